"""Dispatch of recurrent layers: persistent sm_100a kernels (K6) on GPU, plain torch on CPU."""
from __future__ import annotations

import torch


def run(rnn: torch.nn.Module, x, state=None):
    """Returns (outputs [B, T, H], final_state).  CUDA inputs of a covered module (unidirectional,
    batch-first, hidden <= 128) run on ``csrc/ops/rnn.cu``; a missing native library is an error there,
    never a silent fallback.  CPU inputs and uncovered modules use torch."""
    if x.is_cuda:
        from .lstm_persistent import maybe_run

        r = maybe_run(rnn, x, state)
        if r is not None:
            return r
    return rnn(x, state)
