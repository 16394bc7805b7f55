"""CPU restatement of the optimizer update the reference applies to the ArcFace weight.

TEST INFRASTRUCTURE ONLY -- never imported by the product path (only tests/, __graft_entry__.smoke()
and bench.py's CPU legs may import anything under oracle/).

What it restates
    The reference builds ``torch.optim.SGD(..., lr=LR, weight_decay=5e-4, momentum=0.9)``
    (ARCFACE/arc_main.py:252-253) or ``torch.optim.Adam(..., lr=LR)`` (ARCFACE/arc_main.py:249-250) over the
    backbone's and the head's parameters and calls ``optimizer.step()`` once per batch (arc_main.py:330).
    The arithmetic lives in the third-party dependency PyTorch (the reference pins no version; the effective pin
    is the torch in this image, 2.11.0): ``torch/optim/sgd.py::_single_tensor_sgd`` and
    ``torch/optim/adam.py::_single_tensor_adam``.  Their published update rules are restated below in numpy
    fp64; ``tests/test_oracle_optim.py`` pins these functions against ``torch.optim`` itself (CPU) and against
    fixtures under tests/golden/ generated from ``torch.optim`` by tests/golden/make_golden_optim.py.

    ``normalize_rows`` is the weight half of ``F.normalize`` at arc_main.py:159 -- the part of the NEXT step the
    fused kernel produces together with the update.
"""
from __future__ import annotations

import numpy as np


def sgd_step(w, g, buf, *, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
    """One torch.optim.SGD step.  ``buf`` is None before the first step.  Returns (w_new, buf_new)."""
    w = np.asarray(w, dtype=np.float64)
    g = np.asarray(g, dtype=np.float64)
    if weight_decay != 0.0:
        g = g + weight_decay * w
    if momentum != 0.0:
        if buf is None:
            buf = g.copy()
        else:
            buf = momentum * np.asarray(buf, dtype=np.float64) + (1.0 - dampening) * g
        g = g + momentum * buf if nesterov else buf
    return w - lr * g, buf


def adam_step(w, g, exp_avg, exp_avg_sq, step, *, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
    """One torch.optim.Adam step (amsgrad / maximize off).  ``step`` counts from 1.
    Returns (w_new, exp_avg_new, exp_avg_sq_new)."""
    w = np.asarray(w, dtype=np.float64)
    g = np.asarray(g, dtype=np.float64)
    m = np.asarray(exp_avg, dtype=np.float64)
    v = np.asarray(exp_avg_sq, dtype=np.float64)
    b1, b2 = betas
    if weight_decay != 0.0:
        g = g + weight_decay * w
    m = m + (1.0 - b1) * (g - m)
    v = b2 * v + (1.0 - b2) * g * g
    bc1 = 1.0 - b1 ** step
    bc2 = 1.0 - b2 ** step
    denom = np.sqrt(v) / np.sqrt(bc2) + eps
    return w - (lr / bc1) * m / denom, m, v


def normalize_rows(w, eps=1e-12):
    """F.normalize(w, dim=1): (unit rows, 1 / max(||row||, eps))."""
    w = np.asarray(w, dtype=np.float64)
    inv = 1.0 / np.maximum(np.sqrt((w * w).sum(axis=1)), eps)
    return w * inv[:, None], inv
