"""Optimizer-step timing on the GPU box: fused HeadSGD / HeadAdam vs torch.optim (+ the normalisation pass
the next forward would otherwise run), and the whole training step of the head with either optimizer.

    python tools/gpu_bench_optim.py [--classes 1000000] [--dim 512] [--batch 1024] [--iters 20]

Prints one JSON line per measurement (CUDA events on the current stream, after warm-up).
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from ddp_classification_pytorch_b200 import ArcMarginProduct, HeadAdam, HeadSGD, _lib  # noqa: E402


def timed(fn, iters, warmup=5):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    beg, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    beg.record()
    for _ in range(iters):
        fn()
    end.record()
    torch.cuda.synchronize()
    return beg.elapsed_time(end) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--classes", type=int, default=1_000_000)
    ap.add_argument("--dim", type=int, default=512)
    ap.add_argument("--batch", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    C, D, B = a.classes, a.dim, a.batch
    _lib.build()
    torch.manual_seed(0)
    x = torch.randn(B, D, device="cuda")
    y = torch.randint(0, C, (B,), device="cuda")
    elems = C * D

    def head_and_opt(kind, fused):
        torch.manual_seed(0)
        h = ArcMarginProduct(D, C).cuda()
        if kind == "sgd":
            kw = dict(lr=0.01, momentum=0.9, weight_decay=5e-4)  # arc_main.py:252-253
            o = HeadSGD(h, **kw) if fused else torch.optim.SGD(h.parameters(), **kw)
        else:
            kw = dict(lr=1e-3)  # arc_main.py:249-250
            o = HeadAdam(h, **kw) if fused else torch.optim.Adam(h.parameters(), **kw)
        return h, o

    for kind, bytes_per_elem in (("sgd", 22), ("adam", 30)):
        res = {}
        for fused in (True, False):
            h, o = head_and_opt(kind, fused)

            def train_step():
                o.zero_grad(set_to_none=True)
                h.forward_loss(x, y).backward()
                o.step()

            train_step()  # creates state, fills the cache
            ms_step = timed(train_step, a.iters)
            ms_opt = timed(o.step, a.iters)  # grad stays in place: the optimizer step alone
            res["fused" if fused else "torch"] = (ms_step, ms_opt)
            del h, o
            torch.cuda.empty_cache()
        f, t = res["fused"], res["torch"]
        print(json.dumps({
            "optimizer": kind, "classes": C, "dim": D, "batch": B,
            "train_step_ms": {"fused": round(f[0], 4), "torch_optim": round(t[0], 4)},
            "optimizer_step_ms": {"fused(+normalise)": round(f[1], 4), "torch_optim(no normalise)": round(t[1], 4)},
            "fused_kernel_GBps": round(elems * bytes_per_elem / (f[1] * 1e-3) / 1e9, 1),
            "algorithmic_bytes_per_weight_element": bytes_per_elem,
            "samples_per_s_with_optimizer": {"fused": round(B / f[0] * 1e3, 1), "torch_optim": round(B / t[0] * 1e3, 1)},
        }), flush=True)


if __name__ == "__main__":
    main()
