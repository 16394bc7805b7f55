#!/usr/bin/env python
"""Benchmark of the ArcFace head hot path (BASELINE.json metric):

    ArcFace head fwd+bwd samples/s, 512-d embeddings x 1M classes, batch 1024

``python bench.py --gpus N --steps K --warmup W`` (N > 1: launched by torchrun, one rank per GPU;
the classes are sharded by class over the N ranks -- total work fixed => "strong" scaling).
A step = normalise + cosine logits + margin + cross-entropy + dX + dW for one batch.
``--config cfg2|cfg3|cfg4`` selects the other shapes BASELINE.json names (default cfg4 = configs[3], the one
the metric is quoted on).

Printed JSON (rank 0, one line):
  value         samples/s with inputs resident in HBM, CUDA-event timed, max over ranks
  e2e           the same through the public nn.Module API with the step's inputs copied from
                pinned host memory and the loss read back, inside the timed region
  roofline      dominant kernel vs the measured peaks (MEASURED_PEAKS.json): burst AND sustained bf16 fractions,
                `frac` taken against the peak of the regime the timed region ran in, HBM fraction beside it
  sustained     a second, long pass (>= 2 s of load) of the same step: the power-capped steady state
  parity        N > 1: the class-sharded result against ONE un-sharded run of the same inputs on rank 0
                N = 1: the fused result against plain fp32 torch on a sample of rows (all classes)
                (outside every timed region; the run exits non-zero when a bound is broken)
  cpu_baseline  the oracle's torch-CPU port of the reference head on a bounded sample
``--impl reference`` times only that CPU port (the reference itself is eager PyTorch whose CUDA
path is pinned to a GPU by arc_main.py:168; its CPU arithmetic is what runs on the host cores).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

# the shapes BASELINE.json names (configs[0] is the CPU-only case: a parity-test case, not a bench line)
CONFIGS = {
    "cfg2": {"B": 256, "D": 2048, "C": 1000, "name": "BASELINE.json configs[1]",
             "metric": "ArcFace head fwd+bwd samples/s (2048-d, 1000 classes, batch 256)"},
    "cfg3": {"B": 512, "D": 512, "C": 100_000, "name": "BASELINE.json configs[2]",
             "metric": "ArcFace head fwd+bwd samples/s (512-d, 100k classes, batch 512)"},
    "cfg4": {"B": 1024, "D": 512, "C": 1_000_000, "name": "BASELINE.json configs[3]",
             "metric": "ArcFace head fwd+bwd samples/s (512-d, 1M classes, batch 1024)"},
}
S_SCALE, MARGIN = 30.0, 0.5
UNIT = "samples/s"
W_BLOCK = 4096  # rows of W that share one generator seed: the matrix is the same for every shard count


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="cfg4", choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=0, help="override the config's batch")
    ap.add_argument("--classes", type=int, default=0, help="override the config's class count")
    ap.add_argument("--dim", type=int, default=0, help="override the config's feature size")
    ap.add_argument("--easy-margin", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--sustained-seconds", type=float, default=2.5,
                    help="length of the second, power-capped pass (0 = skip it)")
    ap.add_argument("--cpu-sample-rows", type=int, default=0,
                    help="rows of the batch the CPU port runs (0 = 256 beside the GPU number, the full batch for --impl reference)")
    ap.add_argument("--kernel-flags", type=lambda v: int(v, 0), default=0, help="ARCB_* flags for A/B runs")
    a = ap.parse_args()
    cfg = CONFIGS[a.config]
    a.named = not (a.batch or a.classes or a.dim)
    a.batch = a.batch or cfg["B"]
    a.classes = a.classes or cfg["C"]
    a.dim = a.dim or cfg["D"]
    a.metric = cfg["metric"] if a.named else f"ArcFace head fwd+bwd samples/s ({a.dim}-d, {a.classes} classes, batch {a.batch})"
    return a


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            d = json.load(fh)
        return {"burst": float(d["bf16_tflops"]), "sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                "hbm": float(d["hbm_gbs"]), "src": "measured (MEASURED_PEAKS.json)"}
    except Exception:
        return {"burst": 1590.0, "sustained": 1400.0, "hbm": 6650.0, "src": "fallback (B200_PROFILING.md)"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def make_weight_rows(lo, hi, D, C, device):
    """Rows [lo, hi) of the synthetic weight (xavier_uniform bound of the full [C, D] matrix, arc_main.py:146).

    The matrix is defined block by block -- rows [k * W_BLOCK, (k + 1) * W_BLOCK) come from a generator seeded with
    1000 + k -- so every shard count sees the SAME matrix and no rank ever has to hold more than its own rows."""
    bound = (6.0 / (C + D)) ** 0.5
    out = torch.empty(hi - lo, D, dtype=torch.float32, device=device)
    gw = torch.Generator(device=device)
    for k in range(lo // W_BLOCK, (hi + W_BLOCK - 1) // W_BLOCK):
        r0, r1 = k * W_BLOCK, min((k + 1) * W_BLOCK, C)
        gw.manual_seed(1000 + k)
        blk = (torch.rand(r1 - r0, D, generator=gw, device=device) * 2 - 1) * bound
        a, b = max(lo, r0), min(hi, r1)
        out[a - lo:b - lo] = blk[a - r0:b - r0]
    return out


def make_batch(B, D, C, seed=0):
    """X ~ N(0,1), labels uniform; identical on every rank and for every shard count."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(B, D, generator=g)
    y = torch.randint(0, C, (B,), generator=g)
    return x, y


def cpu_port_rate(B_rows, D, C, easy, steps, warmup, threads):
    """samples/s of the oracle's torch-CPU port (fwd + autograd bwd) on B_rows rows, same C and D."""
    from oracle.arcface_oracle import TorchPort
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    bound = (6.0 / (C + D)) ** 0.5
    w = (torch.rand(C, D, generator=g) * 2 - 1) * bound
    x = torch.randn(B_rows, D, generator=g)
    y = torch.randint(0, C, (B_rows,), generator=g)
    port = TorchPort(w, S_SCALE, MARGIN, easy)
    for _ in range(warmup):
        port.step(x, y)
    t0 = time.perf_counter()
    for _ in range(steps):
        port.step(x, y)
    dt = (time.perf_counter() - t0) / max(1, steps)
    return B_rows / dt, dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU arm runs once, on rank 0
    threads = os.cpu_count() or 1
    # always the same sample: the FULL batch of the config (cfg4: ~10 s per step on 16 cores), unless the caller fixes
    # another row count -- never a size probed at run time (the per-sample cost of a small sample is dominated by the
    # weight-side passes and is not comparable)
    rows = args.cpu_sample_rows or args.batch
    rate, dt = cpu_port_rate(rows, args.dim, args.classes, args.easy_margin, args.steps, args.warmup, threads)
    sample = (f"{rows} of {args.batch} batch rows per step, full C={args.classes}, D={args.dim}; torch-CPU fp32 port of "
              f"the reference head (oracle.TorchPort), {dt:.2f} s/step, {args.steps} timed steps after {args.warmup} warm-up")
    line = {
        "impl": "reference", "metric": args.metric, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    label = CONFIGS[args.config]["name"] if args.named else "custom shape"
    big = args.classes // world * args.dim * 4 > 126e6
    return {
        "workload": f"ArcFace head (ArcMarginProduct + CrossEntropyLoss) fwd+bwd, B={args.batch}, D={args.dim}, "
                    f"C={args.classes}, s=30, m=0.5, easy_margin={bool(args.easy_margin)} ({label})",
        "batch": args.batch, "dim": args.dim, "classes": args.classes,
        "parallelism": f"class-sharded x{world}" if world > 1 else "single GPU",
        "cache": ("inputs larger than L2: fp32 W shard read every step "
                  f"({args.classes // world * args.dim * 4 / 1e6:.0f} MB/GPU vs 126 MB L2)") if big else
                 "L2 flushed between timed steps (a 256 MB buffer is overwritten; the fills are timed on their own and subtracted)",
    }


def cosine(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a @ b) / (a.norm() * b.norm() + 1e-300))


def sample_reference_fp32(x, y, w, rows, s, m, easy):
    """Plain fp32 torch (the reference's op sequence, arc_main.py:159-172 + CrossEntropyLoss) for `rows` of the
    batch against ALL classes; returns per-row lse, target logit, and d(sum of those rows' losses)/dx."""
    import math
    F = torch.nn.functional
    xs = x[rows].detach().clone().requires_grad_(True)
    cos = F.linear(F.normalize(xs), F.normalize(w))
    sine = torch.sqrt((1.0 - torch.pow(cos, 2)).clamp(0, 1))
    phi = cos * math.cos(m) - sine * math.sin(m)
    if easy:
        phi = torch.where(cos > 0, phi, cos)
    else:
        phi = torch.where(cos > math.cos(math.pi - m), phi, cos - math.sin(math.pi - m) * m)
    hot = torch.zeros_like(cos)
    hot.scatter_(1, y[rows].view(-1, 1), 1)
    out = (hot * phi + (1.0 - hot) * cos) * s
    lse = torch.logsumexp(out, dim=1)
    zt = out.gather(1, y[rows].view(-1, 1)).squeeze(1)
    (lse - zt).sum().backward()
    top = out.detach().topk(min(4, out.shape[1]), 1, True, True)  # arc_main.py:342 (one more for the 3rd/4th gap)
    return lse.detach(), zt.detach(), xs.grad.detach(), top.values, top.indices


def run_b200(args):
    from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    group = None
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
        group = dist.group.WORLD
    L = _lib.lib()  # raises if the CUDA extension is missing: no fallback

    B, D, C = args.batch, args.dim, args.classes
    head = ArcMarginProduct(D, C, s=S_SCALE, m=MARGIN, easy_margin=args.easy_margin, process_group=group,
                            kernel_flags=args.kernel_flags)
    x_cpu, y_cpu = make_batch(B, D, C)
    head = head.to(dev)
    with torch.no_grad():
        head.weight.copy_(make_weight_rows(head.class_lo, head.class_hi, D, C, dev))
    x_pin, y_pin = x_cpu.pin_memory(), y_cpu.pin_memory()
    x_dev = x_pin.to(dev).requires_grad_(True)
    y_dev = y_pin.to(dev)

    # Timing rule: inputs larger than L2, or L2 flushed between timed iterations.  The fp32 weight shard is re-read
    # from HBM every step; where it is smaller than L2 (cfg2, cfg3 on many GPUs) a 256 MB buffer is overwritten
    # between steps and the time of those fills (measured on their own) is subtracted.
    shard_bytes = (head.class_hi - head.class_lo) * D * 4
    flush = shard_bytes <= 126e6
    flush_buf = torch.empty(256 << 20, dtype=torch.uint8, device=dev) if flush else None

    def step_resident():
        head.weight.grad = None
        x_dev.grad = None
        if flush:
            flush_buf.fill_(1)
        loss = head.forward_loss(x_dev, y_dev)
        loss.backward()
        return loss

    loss_host = torch.empty((), dtype=torch.float32).pin_memory()

    # End-to-end step: this step's inputs come from pinned host memory, the loss goes back to the host.  The copy of
    # step i + 1's inputs runs on a second stream while step i computes (two device buffers); every copy and every
    # read-back is inside the timed region, none is skipped.
    copy_stream = torch.cuda.Stream(device=dev)
    slots = [{"x": torch.empty(B, D, device=dev), "y": torch.empty(B, dtype=torch.int64, device=dev),
              "ready": torch.cuda.Event(), "free": torch.cuda.Event()} for _ in range(2)]
    e2e_state = {"i": 0, "primed": False}

    def enqueue_copy(slot):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(slot["free"])  # the step that last used this buffer has finished with it
            slot["x"].copy_(x_pin, non_blocking=True)
            slot["y"].copy_(y_pin, non_blocking=True)
            slot["ready"].record(copy_stream)

    def step_e2e():
        main = torch.cuda.current_stream(dev)
        if not e2e_state["primed"]:
            for sl in slots:
                sl["free"].record(main)
            enqueue_copy(slots[0])
            e2e_state["primed"] = True
        i = e2e_state["i"]
        cur, nxt = slots[i % 2], slots[(i + 1) % 2]
        enqueue_copy(nxt)                     # next step's inputs: overlaps this step's kernels
        main.wait_event(cur["ready"])
        head.weight.grad = None
        if flush:
            flush_buf.fill_(1)
        xd = cur["x"].detach().requires_grad_(True)
        loss = head.forward_loss(xd, cur["y"])
        loss.backward()
        cur["free"].record(main)
        loss_host.copy_(loss.detach(), non_blocking=True)
        e2e_state["i"] = i + 1
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(step_fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        t0 = time.perf_counter()
        e0.record()
        for _ in range(steps):
            step_fn()
        e1.record()
        host = time.perf_counter() - t0  # the host has ISSUED everything (it runs ahead of the device)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1), wall * 1e3, host * 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        timed.host_ms = ms[2].item()
        return ms[0].item(), ms[1].item()

    def flush_cost_ms(steps):
        if not flush:
            return 0.0
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            flush_buf.fill_(1)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1)

    warm = max(3, args.warmup)  # never fewer than three untimed steps
    for _ in range(warm):
        step_resident()
    _lib.check(L.arcb_check_watchdog(torch.cuda.current_stream().cuda_stream))

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = L.arcb_launch_count()
    dev_ms, _ = timed(step_resident, args.steps)
    host_issue_ms = timed.host_ms / args.steps
    launches = L.arcb_launch_count() - n0
    clocks = sampler.stop() if rank == 0 else None
    fl_ms = flush_cost_ms(args.steps)
    dev_ms_net = max(dev_ms - fl_ms, 1e-6)
    final_loss = step_resident().item()

    # per-kernel timing pass (CUDA events around every launch, on the launching stream): the same K steps again.
    # (a pause before each further pass: every pass should start from the same power / clock state instead of the
    #  later ones inheriting the power-capped clocks of the earlier ones)
    time.sleep(1.0)
    _lib.check(L.arcb_profile_enable(1))
    prof_steps = min(args.steps, 500)  # the record pool holds 8192 launches
    barrier()
    for _ in range(prof_steps):
        step_resident()
    prof = _lib.profile_collect()
    _lib.check(L.arcb_profile_enable(0))

    # end-to-end through the public API with host buffers
    time.sleep(1.0)
    for _ in range(2):
        step_e2e()
    _, e2e_wall_ms = timed(step_e2e, args.steps)
    e2e_wall_net = max(e2e_wall_ms - fl_ms, 1e-6)

    # second figure: the same resident step for >= sustained-seconds of continuous load (power-capped steady state)
    sustained = None
    if args.sustained_seconds > 0:
        time.sleep(1.0)
        long_steps = max(args.steps, int(args.sustained_seconds * 1e3 / max(dev_ms / args.steps, 1e-3)) + 1)
        long_steps = min(long_steps, 20000)
        samp2 = ClockSampler(local)
        if rank == 0:
            samp2.start()
        long_ms, _ = timed(step_resident, long_steps)
        clocks2 = samp2.stop() if rank == 0 else None
        long_net = max(long_ms - flush_cost_ms(long_steps), 1e-6)
        sustained = {"steps": long_steps, "seconds": long_ms * 1e-3, "ms_per_step": long_net / long_steps,
                     "value": B * long_steps / (long_net * 1e-3), "unit": UNIT, "clocks": clocks2}

    ms_per_step = dev_ms_net / args.steps
    value = B * args.steps / (dev_ms_net * 1e-3)
    e2e_value = B * args.steps / (e2e_wall_net * 1e-3)

    # ---- parity, outside every timed region --------------------------------------------------------------------
    parity, parity_ok = None, True
    if not args.no_parity:
        head.weight.grad = None
        x_dev.grad = None
        loss_s = head.forward_loss(x_dev, y_dev, want_top3=True)
        loss_s.backward()
        torch.cuda.synchronize()
        if world > 1:
            # the class-sharded result against ONE un-sharded run of the same inputs on rank 0 (full W: fits one GPU)
            ok_t = torch.ones(1, device=dev)
            if rank == 0:
                full = ArcMarginProduct(D, C, s=S_SCALE, m=MARGIN, easy_margin=args.easy_margin).to(dev)
                with torch.no_grad():
                    full.weight.copy_(make_weight_rows(0, C, D, C, dev))
                xu = x_dev.detach().clone().requires_grad_(True)
                loss_u = full.forward_loss(xu, y_dev, want_top3=True)
                loss_u.backward()
                torch.cuda.synchronize()
                lo, hi = head.class_lo, head.class_hi
                parity = {
                    "against": "un-sharded forward_loss + backward of the same inputs on rank 0",
                    "loss_rel": abs(loss_s.item() - loss_u.item()) / abs(loss_u.item()),
                    "cos_dx": cosine(x_dev.grad, xu.grad),
                    "cos_dw_shard0": cosine(head.weight.grad, full.weight.grad[lo:hi]),
                    "dw_shard0_norm_ratio": float(head.weight.grad.norm() / full.weight.grad[lo:hi].norm()),
                    "top3_equal_frac": float((head.aux["top3"] == full.aux["top3"]).all(dim=1).float().mean()),
                    "lse_max_abs_diff": float((head.aux["lse"] - full.aux["lse"]).abs().max()),
                }
                parity_ok = (parity["loss_rel"] <= 1e-3 and parity["cos_dx"] >= 0.999 and parity["cos_dw_shard0"] >= 0.999
                             and parity["top3_equal_frac"] >= 0.99)
                ok_t.fill_(1.0 if parity_ok else 0.0)
                del full, xu, loss_u
            dist.broadcast(ok_t, src=0)
            parity_ok = bool(ok_t.item() > 0.5)
        else:
            # one GPU: plain fp32 torch on a sample of rows against ALL classes (lse, target logit, dX of those rows)
            n_rows = min(B, 64)
            rows = torch.arange(0, B, max(1, B // n_rows), device=dev)[:n_rows]
            lse_r, zt_r, dx_r, top_v, top_i = sample_reference_fp32(x_dev.detach(), y_dev, head.weight.detach(), rows, S_SCALE,
                                                                    MARGIN, args.easy_margin)
            # argmax / top-3: identical wherever the fp32 reference itself is unambiguous (neighbouring logits further
            # apart than fp32 summation-order noise, 2e-5 * s); the excluded fraction is reported
            k3 = min(3, top_i.shape[1])
            gaps = top_v[:, :-1] - top_v[:, 1:]
            clear = (gaps > 2e-5 * S_SCALE).all(dim=1) if gaps.numel() else torch.ones(len(rows), dtype=torch.bool, device=dev)
            got3 = head.aux["top3"][rows][:, :k3]
            same3 = (got3 == top_i[:, :k3]).all(dim=1)
            lse_g, zt_g = head.aux["lse"][rows], head.aux["target_logit"][rows]
            loss_rows_ref = float((lse_r - zt_r).double().mean())
            loss_rows_got = float((lse_g - zt_g).double().mean())
            parity = {
                "against": f"plain fp32 torch (reference op sequence) on {n_rows} of {B} rows x all {C} classes",
                "loss_rel": abs(loss_rows_got - loss_rows_ref) / abs(loss_rows_ref),
                "lse_max_abs_diff": float((lse_g - lse_r).abs().max()),
                "target_logit_max_abs_diff": float((zt_g - zt_r).abs().max()),
                "cos_dx_rows": cosine(x_dev.grad[rows] * B, dx_r),  # the module's loss is the MEAN over B rows
                "dx_norm_ratio": float((x_dev.grad[rows] * B).norm() / dx_r.norm()),
                "argmax_equal_frac": float((got3[:, 0] == top_i[:, 0]).float().mean()),
                "top3_equal_frac": float(same3.float().mean()),
                "rows_unambiguous_frac": float(clear.float().mean()),
                "top3_equal_on_unambiguous_rows": bool(same3[clear].all()) if bool(clear.any()) else True,
            }
            parity_ok = (parity["loss_rel"] <= 1e-3 and parity["cos_dx_rows"] >= 0.999
                         and parity["target_logit_max_abs_diff"] <= 1e-3 and parity["top3_equal_on_unambiguous_rows"])

    if rank == 0:
        peaks = measured_peaks()
        Cs = head.class_hi - head.class_lo
        flops_per_gemm = 2.0 * B * Cs * D  # per GPU, per step, for each of the three algorithmic GEMMs
        kinds = {k: (ms / prof_steps, n // prof_steps) for k, (ms, n) in prof.items() if n}
        # algorithmic GEMMs per launch kind: the persistent backward does all three backward GEMMs
        # (recompute + dX + dW; only dX and dW count as algorithmic flops)
        gemms_in = {"fwd_gemm": 1, "bwd_g_gemm": 0, "dx_gemm": 1, "dw_gemm": 1, "bwd_persistent": 2}
        # algorithmic HBM bytes per weight element and launch kind (14 B x C x D per step in all):
        # forward = read fp32 W 4 + write bf16 w_hat 2 + read it 2 (the first two only when the kernel normalises);
        # backward = read w_hat 2 + write fp32 dW 4
        fused_norm = "l2norm_w" not in kinds
        bytes_in = {"fwd_gemm": 8.0 if fused_norm else 2.0, "bwd_persistent": 6.0, "dw_gemm": 4.0, "dx_gemm": 2.0,
                    "bwd_g_gemm": 2.0, "l2norm_w": 6.0}
        gemm_kinds = {k: v for k, v in kinds.items() if k in gemms_in}
        dom = max(gemm_kinds, key=lambda k: gemm_kinds[k][0])
        dom_ms, dom_n = gemm_kinds[dom]
        dom_flops = flops_per_gemm * gemms_in[dom]
        dom_bytes = bytes_in.get(dom, 0.0) * Cs * D
        achieved_tf = dom_flops / (dom_ms * 1e-3) / 1e12
        achieved_gbs = dom_bytes / (dom_ms * 1e-3) / 1e9
        # which roof binds this kernel: the larger of its ideal times (tensor at the burst peak, HBM at the copy peak)
        t_tensor, t_hbm = dom_flops / (peaks["burst"] * 1e12), dom_bytes / (peaks["hbm"] * 1e9)
        bound = "tensor" if t_tensor >= t_hbm else "hbm"
        timed_seconds = dev_ms * 1e-3
        regime = "burst" if timed_seconds < 1.0 else "sustained"  # < 1 s of load: clocks have not settled under the cap
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")
        # the ncu capture under profiles/ is of the default workload on one GPU: no figure for anything else
        profiled = (B, C, D, world) == (CONFIGS["cfg4"]["B"], CONFIGS["cfg4"]["C"], CONFIGS["cfg4"]["D"], 1)
        if profiled and os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get(dom)
            except Exception:
                traffic = None
        if bound == "tensor":
            roof = {"bound": "tensor", "achieved": achieved_tf, "peak": peaks[regime], "unit": "TFLOP/s",
                    "frac": achieved_tf / peaks[regime]}
        else:
            roof = {"bound": "hbm", "achieved": achieved_gbs, "peak": peaks["hbm"], "unit": "GB/s",
                    "frac": achieved_gbs / peaks["hbm"]}
        roof.update({
            "kernel": dom, "traffic": traffic, "regime": f"{regime} (timed region {timed_seconds:.3f} s)",
            "peak_source": peaks["src"],
            "frac_burst": achieved_tf / peaks["burst"], "frac_sustained": achieved_tf / peaks["sustained"],
            "frac_hbm": achieved_gbs / peaks["hbm"], "achieved_tflops": achieved_tf, "achieved_gbs": achieved_gbs,
            "launches_per_step": dom_n, "ms_per_step": dom_ms,
            "algorithmic_flops_per_step": dom_flops, "algorithmic_bytes_per_step": dom_bytes,
            "executed_flops_per_step": flops_per_gemm * {"bwd_persistent": 3, "bwd_g_gemm": 1}.get(dom, gemms_in[dom]),
        })
        step_tf = 6.0 * B * C * D / (ms_per_step * 1e-3) / 1e12
        step_gbs = 14.0 * C * D / (ms_per_step * 1e-3) / 1e9
        line = {
            "metric": args.metric, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "warmup_steps_run": warm,
            "ms_per_step": ms_per_step, "host_issue_ms_per_step": host_issue_ms,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "config": workload_config(args, world),
            "loss": final_loss,
            "step_tflops_algorithmic": step_tf,
            "step_frac_of_burst_peak": step_tf / (peaks["burst"] * world),
            "step_frac_of_sustained_peak": step_tf / (peaks["sustained"] * world),
            "step_frac_of_hbm_peak": step_gbs / (peaks["hbm"] * world),
            "roofline": roof,
            "kernels_ms_per_step": {k: round(v[0], 4) for k, v in kinds.items()},
            "kernels_launches_per_step": {k: v[1] for k, v in kinds.items()},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * D * 4 + B * 8, "d2h_bytes_per_step": 4,
                    "ms_per_step": e2e_wall_net / args.steps, "timer": "wall clock between device syncs",
                    "note": "value and e2e are separate passes (1 s apart) in slightly different power states: "
                            "differences below ~3 % between them are not signal"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "parity": parity,
        }
        if flush:
            line["l2_flush"] = {"bytes": 256 << 20, "ms_per_step_subtracted": fl_ms / args.steps}
        if sustained is not None:
            s_tf = 6.0 * B * C * D / (sustained["ms_per_step"] * 1e-3) / 1e12
            sustained["step_frac_of_sustained_peak"] = s_tf / (peaks["sustained"] * world)
            sustained["step_frac_of_burst_peak"] = s_tf / (peaks["burst"] * world)
            line["sustained"] = sustained
        if world == 1 and not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            rows = args.cpu_sample_rows or min(B, 256)  # ~10-20 s of CPU work in total
            try:
                rate, dt = cpu_port_rate(rows, D, C, args.easy_margin, 2, 1, threads)
                line["cpu_baseline"] = {
                    "value": rate, "unit": UNIT, "cores": threads, "kind": "port",
                    "sample": f"{rows} of {B} batch rows per step (2 timed steps after 1 warm-up), full C={C}, D={D}; "
                              f"torch-CPU fp32 port of the reference head (oracle.TorchPort), {dt:.2f} s/step"}
            except Exception as ex:  # e.g. host RAM too small for the 1M-class sample
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": threads, "kind": "port",
                                        "sample": f"failed: {type(ex).__name__}: {ex}"}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if not parity_ok:
        print("bench.py: PARITY CHECK FAILED", file=sys.stderr, flush=True)
        sys.exit(3)


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
