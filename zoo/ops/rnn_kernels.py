"""Dispatch of recurrent layers to the persistent sm_100a kernels (K6) or to cuDNN via torch."""
from __future__ import annotations

import torch


def run(rnn: torch.nn.Module, x, state=None):
    """Returns (outputs [B, T, H], final_state).  The hand-written persistent kernel handles
    single-direction LSTMs at inference (see csrc/ops/rnn.cu); training uses torch autograd."""
    try:
        from .lstm_persistent import maybe_run

        r = maybe_run(rnn, x, state)
        if r is not None:
            return r
    except ImportError:
        pass
    return rnn(x, state)
