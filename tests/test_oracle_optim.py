"""Pins oracle/optim_oracle.py: against the committed torch.optim fixtures and against torch.optim run live."""
import ast
import glob
import os

import numpy as np
import pytest
import torch

from oracle import optim_oracle as oo

OPT_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "optim")
CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(OPT_DIR, "*.npz")))


def load_case(name):
    with np.load(os.path.join(OPT_DIR, name + ".npz")) as z:
        d = {k: z[k] for k in z.files}
    d["kind"] = str(d["kind"])
    d["kwargs"] = ast.literal_eval(str(d["kwargs"]))
    return d


def run_oracle(d):
    """Yield (w, s0, s1) after every step, each step restarted from the fixture's fp32 state (no drift)."""
    kw = d["kwargs"]
    steps = d["grads"].shape[0]
    for t in range(steps):
        w_prev = d["w0"] if t == 0 else d["w"][t - 1]
        if d["kind"] == "sgd":
            buf = None if t == 0 else d["s0"][t - 1]
            w, buf = oo.sgd_step(w_prev, d["grads"][t], buf, **kw)
            yield w, buf, None
        else:
            m = np.zeros_like(d["w0"]) if t == 0 else d["s0"][t - 1]
            v = np.zeros_like(d["w0"]) if t == 0 else d["s1"][t - 1]
            yield oo.adam_step(w_prev, d["grads"][t], m, v, t + 1, **kw)


def test_fixtures_present():
    assert len(CASES) >= 6


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_torch_optim_fixture(name):
    d = load_case(name)
    for t, (w, s0, s1) in enumerate(run_oracle(d)):
        np.testing.assert_allclose(w, d["w"][t], rtol=2e-6, atol=2e-8)  # torch computes in fp32
        if s0 is not None:
            np.testing.assert_allclose(s0, d["s0"][t], rtol=2e-6, atol=3e-8)  # fp32 cancellation near 0
        if s1 is not None:
            np.testing.assert_allclose(s1, d["s1"][t], rtol=2e-6, atol=1e-12)


@pytest.mark.parametrize("kind,kw", [
    ("sgd", dict(lr=0.01, momentum=0.9, weight_decay=5e-4)),
    ("sgd", dict(lr=0.3, momentum=0.5, nesterov=True)),
    ("adam", dict(lr=1e-3)),
    ("adam", dict(lr=1e-2, betas=(0.5, 0.9), weight_decay=0.1)),
])
def test_oracle_matches_torch_optim_live_fp64(kind, kw):
    """Multi-step trajectory in fp64 on both sides: agreement to rounding."""
    g = torch.Generator().manual_seed(7)
    w = torch.nn.Parameter(torch.randn(9, 20, generator=g, dtype=torch.float64))
    opt = (torch.optim.SGD if kind == "sgd" else torch.optim.Adam)([w], foreach=False, **kw)
    wn, s0, s1 = w.detach().numpy().copy(), None, None
    if kind == "adam":
        s0, s1 = np.zeros_like(wn), np.zeros_like(wn)
    for t in range(6):
        gr = torch.randn(9, 20, generator=g, dtype=torch.float64)
        w.grad = gr.clone()
        opt.step()
        if kind == "sgd":
            wn, s0 = oo.sgd_step(wn, gr.numpy(), s0, **kw)
        else:
            wn, s0, s1 = oo.adam_step(wn, gr.numpy(), s0, s1, t + 1, **kw)
        np.testing.assert_allclose(wn, w.detach().numpy(), rtol=1e-12, atol=1e-14)


def test_normalize_rows_matches_torch():
    g = torch.Generator().manual_seed(3)
    w = torch.randn(11, 33, generator=g, dtype=torch.float64)
    w[4] = 0  # clamped row: F.normalize divides by eps
    what, inv = oo.normalize_rows(w.numpy())
    np.testing.assert_allclose(what, torch.nn.functional.normalize(w).numpy(), rtol=1e-13, atol=0)
    assert inv[4] == 1e12
