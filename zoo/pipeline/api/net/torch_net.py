"""TorchNet: a PyTorch module as a zoo model.

Reference: pyzoo/zoo/pipeline/api/net/torch_net.py:36-121 and zoo/.../pipeline/api/net/TorchNet.scala:39-241 --
`from_pytorch(module, input)` traces the module to TorchScript, ships the bytes to every executor, and each
step copies the flat weight vector into libtorch, runs forward/backward through JNI and copies the flat
gradient back (three full-model copies per iteration).  Here the module IS the model: `TorchNet` is a thin
`KerasNet` around it (so `compile / fit / evaluate / predict`, NNFrames and `Estimator` accept it), its
parameters are re-pointed into the flat arena once, and nothing is copied per step.  `from_pytorch` still
traces when asked (`TorchNet.from_pytorch(module, input)` -> `.traced`) and `savePytorch` writes the
TorchScript archive the reference writes.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn

from ..keras.engine.topology import KerasNet


class TorchNet(KerasNet):
    def __init__(self, module_or_path, bigdl_type="float"):
        super().__init__()
        if isinstance(module_or_path, nn.Module):
            self.module = module_or_path
        else:
            self.module = torch.jit.load(module_or_path, map_location="cpu")
        self.traced = None

    @staticmethod
    def get_sample_input(input):
        """list/tuple of ints = one input shape; list of lists = several; tensors are used as they are."""
        if isinstance(input, torch.Tensor):
            return input
        if isinstance(input, (list, tuple)) and input and all(isinstance(i, int) for i in input):
            return torch.rand(*input)
        if isinstance(input, (list, tuple)):
            return tuple(TorchNet.get_sample_input(i) for i in input)
        raise TypeError(f"cannot build a sample input from {type(input)}")

    @staticmethod
    def from_pytorch(module, input=None, check_trace=True):
        net = TorchNet(module)
        if input is not None:
            sample = TorchNet.get_sample_input(input)
            was = module.training
            module.eval()
            try:
                net.traced = torch.jit.trace(module, sample, check_trace=check_trace)
            finally:
                module.train(was)
        return net

    def forward(self, *xs):
        return self.module(*xs)

    def savePytorch(self, path):
        m = self.traced if self.traced is not None else torch.jit.script(self.module)
        torch.jit.save(m, path)

    save_pytorch = savePytorch

    def predict(self, x, batch_per_thread=1, distributed=True, batch_size=None):
        return super().predict(x, batch_per_thread=batch_per_thread, batch_size=batch_size or max(1, batch_per_thread) * 256)

    def get_weights(self):
        """Flat weight vector, like the reference's `getWeightNative`."""
        return torch.cat([p.detach().reshape(-1) for p in self.module.parameters()]).cpu().numpy()

    def set_weights(self, flat):
        flat = torch.as_tensor(np.asarray(flat), dtype=torch.float32)
        off = 0
        with torch.no_grad():
            for p in self.module.parameters():
                n = p.numel()
                p.copy_(flat[off:off + n].reshape(p.shape))
                off += n
        if off != flat.numel():
            raise ValueError(f"expected {off} weights, got {flat.numel()}")
        return self
