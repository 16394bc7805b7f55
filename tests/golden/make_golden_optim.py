"""Generate tests/golden/optim/*.npz from torch.optim ITSELF (CPU, the torch of this image).

    python tests/golden/make_golden_optim.py

The reference's optimizer step (ARCFACE/arc_main.py:249-253, :330) is plain ``torch.optim.SGD`` /
``torch.optim.Adam``; each fixture stores a seeded weight, a sequence of gradients and the weight / optimizer
state torch.optim produces after every step.  They pin oracle/optim_oracle.py (tests/test_oracle_optim.py) and
the CUDA kernels (tests/test_gpu_optim.py).
"""
from __future__ import annotations

import os

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "optim")

# name, kind, rows, D, steps, kwargs
CASES = [
    ("sgd_trainer", "sgd", 37, 72, 4, dict(lr=0.01, momentum=0.9, weight_decay=5e-4)),       # arc_main.py:252-253
    ("sgd_plain", "sgd", 19, 64, 3, dict(lr=0.1)),
    ("sgd_nesterov", "sgd", 23, 70, 3, dict(lr=0.05, momentum=0.8, weight_decay=1e-3, nesterov=True)),
    ("sgd_dampening", "sgd", 16, 256, 3, dict(lr=0.02, momentum=0.7, dampening=0.3)),
    ("adam_trainer", "adam", 37, 72, 4, dict(lr=1e-3)),                                       # arc_main.py:249-250
    ("adam_wd", "adam", 21, 70, 3, dict(lr=3e-3, betas=(0.8, 0.95), eps=1e-6, weight_decay=1e-2)),
]


def main():
    os.makedirs(OUT, exist_ok=True)
    for name, kind, rows, D, steps, kw in CASES:
        g = torch.Generator().manual_seed(sum(map(ord, name)))
        bound = (6.0 / (rows + D)) ** 0.5
        w = torch.nn.Parameter((torch.rand(rows, D, generator=g) * 2 - 1) * bound)
        grads = torch.randn(steps, rows, D, generator=g) * 0.05
        opt = (torch.optim.SGD if kind == "sgd" else torch.optim.Adam)([w], foreach=False, **kw)
        out = {"w0": w.detach().numpy().copy(), "grads": grads.numpy()}
        ws, s0, s1 = [], [], []
        for t in range(steps):
            w.grad = grads[t].clone()
            opt.step()
            ws.append(w.detach().numpy().copy())
            st = opt.state[w]
            if kind == "sgd":
                buf = st.get("momentum_buffer")
                s0.append(np.zeros((rows, D), np.float32) if buf is None else buf.numpy().copy())
            else:
                s0.append(st["exp_avg"].numpy().copy())
                s1.append(st["exp_avg_sq"].numpy().copy())
        out["w"] = np.stack(ws)
        out["s0"] = np.stack(s0)
        if s1:
            out["s1"] = np.stack(s1)
        out["kind"] = np.array(kind)
        out["kwargs"] = np.array(repr(kw))
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
        print(name, rows, D, steps, kw)


if __name__ == "__main__":
    main()
