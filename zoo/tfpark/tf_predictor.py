"""`zoo.tfpark.tf_predictor.TFPredictor` (reference: pyzoo/zoo/tfpark/tf_predictor.py:27-90).

The reference freezes (session, inputs, outputs) into a TFNet and maps it over the TFDataset's prediction data.  The same two
steps here: the model is a `TFNet` (a frozen GraphDef imported into torch, `zoo.tfpark.graphdef`), a zoo Keras-API model, a
`tfpark.KerasModel` or any torch module, and `predict()` runs it over `dataset.get_prediction_data()` in `batch_per_thread`
batches on the GPU, returning one array per record.
"""
from __future__ import annotations

import numpy as np
import torch

from .tf_dataset import TFDataset


class TFPredictor:
    def __init__(self, sess, outputs=None, inputs=None, dataset=None):
        """`sess`: the model (see module docstring) -- the slot the reference uses for the tf.Session; `outputs`: optional list of
        output indices to keep when the model returns several tensors."""
        import torch.nn as nn

        from .model import KerasModel

        model = sess.model if isinstance(sess, KerasModel) else sess
        if not isinstance(model, nn.Module):
            raise NotImplementedError("TFPredictor(sess, outputs) freezes a live tf.Session, and there is no TensorFlow runtime in zoo-b200; "
                                      "pass a TFNet (TFNet.from_export_folder / from_frozen_graph), a Keras-API model or a torch module")
        if not isinstance(dataset, TFDataset):
            raise ValueError("dataset must be a TFDataset")
        if dataset.batch_per_thread <= 0:
            raise ValueError("You should set batch_per_thread on TFDataset instead of batch_size for prediction")
        self.model, self.outputs, self.inputs, self.dataset = model, outputs, inputs, dataset

    @classmethod
    def from_outputs(cls, sess, outputs=None, dataset=None):
        return cls(sess, outputs, None, dataset)

    @classmethod
    def from_keras(cls, keras_model, dataset):
        return cls(keras_model, None, None, dataset)

    def predict(self):
        dev = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        net = self.model.to(dev).eval()
        outs = []
        with torch.no_grad():
            for mb in self.dataset.get_prediction_data().batches(self.dataset.batch_per_thread, train=False):
                xs = mb.inputs if isinstance(mb.inputs, (tuple, list)) else (mb.inputs,)
                o = net(*[t.to(dev) for t in xs])
                if isinstance(o, (tuple, list)):
                    o = [o[i] for i in self.outputs] if self.outputs else list(o)
                    cols = [t.float().cpu().numpy() for t in o]
                    outs.extend([c[i] for c in cols] for i in range(len(cols[0])))
                else:
                    outs.extend(np.asarray(o.float().cpu()))
        return outs
