"""Quick A/B timer for the head's kernels: forward and backward timed separately with CUDA events.

    [ARCB_* tuning env vars] python tools/gpu_time_bwd.py [--classes N] [--batch B] [--dim D] [--iters K] [--flags 0x..] [--tag name]

Prints one JSON line; meant to be called several times inside one gpurun command with different flags / env vars.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ddp_classification_pytorch_b200 import _lib  # noqa: E402
from ddp_classification_pytorch_b200.head import DEFAULT_CHUNK_CLASSES, CudaEngine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--classes", type=int, default=1_000_000)
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--flags", type=lambda v: int(v, 0), default=0)
    ap.add_argument("--tag", default="")
    a = ap.parse_args()
    B, D, C = a.batch, a.dim, a.classes
    g = torch.Generator(device="cuda").manual_seed(0)
    x = torch.randn(B, D, generator=g, device="cuda")
    w = (torch.rand(C, D, generator=g, device="cuda") - 0.5) * 0.01
    y = torch.randint(0, C, (B,), generator=g, device="cuda")
    eng = CudaEngine()
    L = _lib.lib()
    cfg = {"s": 30.0, "m": 0.5, "easy_margin": False, "class_offset": 0, "chunk_classes": DEFAULT_CHUNK_CLASSES,
           "flags": a.flags}
    one = torch.ones((), device="cuda")

    def fwd():
        stats, state = eng.shard_forward(x, w, y, cfg)
        lse, zt, loss = eng.merge_loss(stats, 1, B)
        return state, lse, zt, loss

    def bwd(state, lse, zt, loss):
        return eng.shard_backward(state, cfg, lse=lse, zt=zt, grad_loss=one)

    for _ in range(3):
        s = fwd()
        bwd(*s)
    torch.cuda.synchronize()
    L.arcb_profile_enable(1)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3 * a.iters)]
    for i in range(a.iters):
        ev[3 * i].record()
        s = fwd()
        ev[3 * i + 1].record()
        bwd(*s)
        ev[3 * i + 2].record()
    torch.cuda.synchronize()
    prof = _lib.profile_collect()
    L.arcb_profile_enable(0)
    rc = L.arcb_check_watchdog(torch.cuda.current_stream().cuda_stream)
    f = sorted(ev[3 * i].elapsed_time(ev[3 * i + 1]) for i in range(a.iters))
    b = sorted(ev[3 * i + 1].elapsed_time(ev[3 * i + 2]) for i in range(a.iters))
    env = {k: v for k, v in os.environ.items() if k.startswith("ARCB_")}
    kinds = {k: round(ms / a.iters, 4) for k, (ms, n) in prof.items() if n}
    print(json.dumps({"tag": a.tag, "flags": hex(a.flags), "env": env, "shape": [B, D, C], "fwd_ms_med": round(f[len(f) // 2], 4),
                      "fwd_ms_min": round(f[0], 4), "bwd_ms_med": round(b[len(b) // 2], 4), "bwd_ms_min": round(b[0], 4),
                      "loss": float(s[3]), "kernels_ms": kinds, "watchdog": rc}), flush=True)


if __name__ == "__main__":
    main()
