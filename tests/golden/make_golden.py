"""Generate the golden vectors under tests/golden/ from the REFERENCE ITSELF.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

For every case it runs the reference's ``ArcMarginProduct`` (AST-extracted from
/root/reference/ARCFACE/arc_main.py:130-176, executed unchanged on CPU, see oracle/ref_extract.py)
followed by ``nn.CrossEntropyLoss`` and ``loss.backward()`` exactly as arc_main.py:326-329 does,
and stores inputs + outputs as ``<case>.npz``.  The committed fixtures pin the oracle
(tests/test_oracle.py) and, through it and directly, the CUDA path (tests/test_gpu_parity.py).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle.ref_extract import reference_forward_backward  # noqa: E402

# name, B, D, C, s, m, easy_margin, seed, planted fraction
CASES = [
    ("cfg1_hard", 64, 512, 10, 30.0, 0.5, False, 0, 0.25),   # BASELINE.json configs[0]
    ("cfg1_easy", 64, 512, 10, 30.0, 0.5, True, 0, 0.25),
    ("small_hard", 32, 64, 1000, 30.0, 0.5, False, 1, 0.25),
    ("small_easy", 32, 64, 1000, 30.0, 0.5, True, 1, 0.25),
    ("trainer_easy", 32, 256, 2173, 30.0, 0.5, True, 2, 0.2),  # arc_main.py:234 operating point
    ("odd_hard_s64", 40, 72, 333, 64.0, 0.35, False, 3, 0.3),  # ragged sizes, other s/m
]


def make_inputs(B, D, C, seed, planted):
    """Seeded inputs with planted rows that hit every margin branch."""
    g = torch.Generator().manual_seed(seed)
    bound = (6.0 / (C + D)) ** 0.5  # xavier_uniform_ bound (arc_main.py:146)
    w = (torch.rand(C, D, generator=g) * 2 - 1) * bound
    x = torch.randn(B, D, generator=g)
    y = torch.randint(0, C, (B,), generator=g)
    n_plant = int(B * planted)
    for r in range(n_plant):
        wy = w[y[r]]
        wy = wy / wy.norm()
        noise = torch.randn(D, generator=g)
        noise = noise - (noise @ wy) * wy
        noise = noise / noise.norm()
        # target cosines spread over (-1, 1): below th, around 0, near +/-1
        c = [-0.995, -0.93, -0.87, -0.5, -1e-3, 1e-3, 0.3, 0.8, 0.97, 0.9995][r % 10]
        x[r] = (c * wy + (1 - c * c) ** 0.5 * noise) * (0.5 + r)
    return x, w, y


def main():
    torch.manual_seed(0)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    for name, B, D, C, s, m, easy, seed, planted in CASES:
        x, w, y = make_inputs(B, D, C, seed, planted)
        ref = reference_forward_backward(x, w, y, s=s, m=m, easy_margin=easy)
        out = os.path.join(HERE, f"{name}.npz")
        np.savez_compressed(
            out,
            x=x.numpy(), w=w.numpy(), label=y.numpy(),
            s=np.float64(s), m=np.float64(m), easy_margin=np.bool_(easy),
            logits=ref["logits"].numpy(), loss=np.float64(ref["loss"]),
            dx=ref["dx"].numpy(), dw=ref["dw"].numpy(),
            torch_version=np.str_(torch.__version__),
        )
        print(f"{name}: B={B} D={D} C={C} loss={ref['loss']:.6f} -> {os.path.relpath(out)}"
              f" ({os.path.getsize(out) / 1e6:.2f} MB)")

    # Known-answer tests: rows built so the target cosine is exactly +/-1 or 0.
    D, C = 8, 4
    w = torch.eye(C, D)
    x = torch.stack([w[0] * 3.0, -w[1] * 2.0, torch.zeros(D), w[3] + 0.0])
    y = torch.tensor([0, 1, 2, 0])  # cos_t = +1, -1, 0 (zero row), 0 (orthogonal)
    kat = {}
    for easy in (False, True):
        Ref_out = reference_forward_backward(x, w, y, s=30.0, m=0.5, easy_margin=easy)
        kat[f"logits_{'easy' if easy else 'hard'}"] = Ref_out["logits"].numpy()
    np.savez_compressed(os.path.join(HERE, "kat.npz"), x=x.numpy(), w=w.numpy(), label=y.numpy(), **kat)
    print("kat: ", {k: v[[0, 1, 2, 3], [0, 1, 2, 0]] / 30.0 for k, v in kat.items()})


if __name__ == "__main__":
    main()
