#!/usr/bin/env python
"""Benchmark of the ArcFace head hot path (BASELINE.json metric):

    ArcFace head fwd+bwd samples/s, 512-d embeddings x 1M classes, batch 1024

``python bench.py --gpus N --steps K --warmup W`` (N > 1: launched by torchrun, one rank per GPU;
the 1M classes are sharded by class over the N ranks -- total work fixed => "strong" scaling).
A step = normalise + cosine logits + margin + cross-entropy + dX + dW for one batch.

Printed JSON (rank 0, one line):
  value         samples/s with inputs resident in HBM, CUDA-event timed, max over ranks
  e2e           the same through the public nn.Module API with the step's inputs copied from
                pinned host memory and the loss read back, inside the timed region
  roofline      dominant kernel vs the measured bf16 tensor peak (MEASURED_PEAKS.json)
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

# BASELINE.json configs[3]: the configuration the metric is quoted on (fits one B200: W = 2 GB)
WORKLOAD = {"B": 1024, "D": 512, "C": 1_000_000, "s": 30.0, "m": 0.5, "easy_margin": False}
METRIC = "ArcFace head fwd+bwd samples/s (512-d, 1M classes, batch 1024)"
UNIT = "samples/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=WORKLOAD["B"])
    ap.add_argument("--classes", type=int, default=WORKLOAD["C"])
    ap.add_argument("--dim", type=int, default=WORKLOAD["D"])
    ap.add_argument("--easy-margin", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-sample-rows", type=int, default=0, help="rows of the batch the CPU port runs (0 = auto)")
    return ap.parse_args()


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as fh:
            d = json.load(fh)
        return {"burst": float(d["bf16_tflops"]), "sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                "hbm": float(d["hbm_gbs"]), "src": "measured"}
    except Exception:
        return {"burst": 1590.0, "sustained": 1400.0, "hbm": 6650.0, "src": "fallback"}


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


def make_inputs(B, D, C_lo, C_hi, C, device, seed=0):
    """Synthetic inputs of the named shapes: X ~ N(0,1), W ~ xavier_uniform, labels uniform.

    Every rank draws the same X / labels; W rows are generated per shard (seeded by the shard) so
    that no rank ever holds the full matrix.
    """
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randn(B, D, generator=g)
    y = torch.randint(0, C, (B,), generator=g)
    bound = (6.0 / (C + D)) ** 0.5
    gw = torch.Generator(device=device).manual_seed(1000 + C_lo)
    w = (torch.rand(C_hi - C_lo, D, generator=gw, device=device) * 2 - 1) * bound
    return x, y, w


def cpu_port_rate(B_rows, D, C, easy, steps, warmup, threads):
    """samples/s of the oracle's torch-CPU port (fwd + autograd bwd) on B_rows rows, same C and D."""
    from oracle.arcface_oracle import TorchPort
    torch.set_num_threads(threads)
    g = torch.Generator().manual_seed(0)
    bound = (6.0 / (C + D)) ** 0.5
    w = (torch.rand(C, D, generator=g) * 2 - 1) * bound
    x = torch.randn(B_rows, D, generator=g)
    y = torch.randint(0, C, (B_rows,), generator=g)
    port = TorchPort(w, WORKLOAD["s"], WORKLOAD["m"], easy)
    for _ in range(warmup):
        port.step(x, y)
    t0 = time.perf_counter()
    for _ in range(steps):
        port.step(x, y)
    dt = (time.perf_counter() - t0) / max(1, steps)
    return B_rows / dt, dt


def size_cpu_sample(args, threads, seconds_per_step):
    """Rows of the batch the CPU port can process in ~seconds_per_step.

    One CPU step costs a + b * rows (a: normalising / differentiating the full [C, D] weight, b: the B x C
    passes), so two probing steps give the model and the sample is sized from it (never above the batch)."""
    r1, r2 = 16, 48
    _, t1 = cpu_port_rate(r1, args.dim, args.classes, args.easy_margin, 1, 0, threads)
    _, t2 = cpu_port_rate(r2, args.dim, args.classes, args.easy_margin, 1, 0, threads)
    b = max((t2 - t1) / (r2 - r1), 1e-6)
    a = max(t1 - b * r1, 0.0)
    rows = int((seconds_per_step - a) / b) if seconds_per_step > a else r1
    return max(r1, min(args.batch, rows)), a, b


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # the CPU arm runs once, on rank 0
    threads = os.cpu_count() or 1
    if args.cpu_sample_rows:
        rows = args.cpu_sample_rows
    else:  # K + W steps (plus the two probes) should end within ~2 minutes
        rows, _, _ = size_cpu_sample(args, threads, 100.0 / max(1, args.steps + args.warmup))
    rate, dt = cpu_port_rate(rows, args.dim, args.classes, args.easy_margin, args.steps, args.warmup, threads)
    sample = (f"{rows} of {args.batch} batch rows per step, full C={args.classes}, D={args.dim}; torch-CPU fp32 port of "
              f"the reference head (oracle.TorchPort), {dt:.2f} s/step")
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, 1),
        "cpu_baseline": {"value": rate, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def workload_config(args, world):
    return {
        "workload": f"ArcFace head (ArcMarginProduct + CrossEntropyLoss) fwd+bwd, B={args.batch}, D={args.dim}, "
                    f"C={args.classes}, s=30, m=0.5, easy_margin={bool(args.easy_margin)} (BASELINE.json configs[3])",
        "batch": args.batch, "dim": args.dim, "classes": args.classes,
        "parallelism": f"class-sharded x{world}" if world > 1 else "single GPU",
        "cache": "inputs larger than L2: fp32 W shard read every step "
                 f"({args.classes // world * args.dim * 4 / 1e6:.0f} MB/GPU vs 126 MB L2)",
    }


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
    head = ArcMarginProduct(D, C, s=WORKLOAD["s"], m=WORKLOAD["m"], easy_margin=args.easy_margin, process_group=group)
    x_cpu, y_cpu, w = make_inputs(B, D, head.class_lo, head.class_hi, C, dev)
    head = head.to(dev)
    with torch.no_grad():
        head.weight.copy_(w)
    del w
    x_pin, y_pin = x_cpu.pin_memory(), y_cpu.pin_memory()
    x_dev = x_pin.to(dev).requires_grad_(True)
    y_dev = y_pin.to(dev)

    def step_resident():
        head.weight.grad = None
        x_dev.grad = None
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
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1), wall * 1e3], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return ms[0].item(), ms[1].item()

    for _ in range(max(3, args.warmup)):
        step_resident()
    _lib.check(L.arcb_check_watchdog(torch.cuda.current_stream().cuda_stream))

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    n0 = L.arcb_launch_count()
    dev_ms, _ = timed(step_resident, args.steps)
    launches = L.arcb_launch_count() - n0
    clocks = sampler.stop() if rank == 0 else None
    final_loss = step_resident().item()

    # per-kernel timing pass (CUDA events around every launch, on the launching stream): the same K steps again,
    # straight after the timed region so that clocks / power state are the ones the headline number saw
    # (a pause before each further pass: the three passes should start from the same power / clock state instead of
    #  the later ones inheriting the power-capped clocks of the earlier ones)
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

    ms_per_step = dev_ms / args.steps
    value = B * args.steps / (dev_ms * 1e-3)
    e2e_value = B * args.steps / (e2e_wall_ms * 1e-3)

    if rank == 0:
        peaks = measured_peaks()
        Cs = head.class_hi - head.class_lo
        flops_per_gemm = 2.0 * B * Cs * D  # per GPU, per step, for each of the three algorithmic GEMMs
        kinds = {k: (ms / prof_steps, n // prof_steps) for k, (ms, n) in prof.items() if n}
        # algorithmic GEMMs per launch kind: the persistent backward does all three backward GEMMs
        # (recompute + dX + dW; only dX and dW count as algorithmic flops)
        gemms_in = {"fwd_gemm": 1, "bwd_g_gemm": 0, "dx_gemm": 1, "dw_gemm": 1, "bwd_persistent": 2}
        gemm_kinds = {k: v for k, v in kinds.items() if k in gemms_in}
        dom = max(gemm_kinds, key=lambda k: gemm_kinds[k][0])
        dom_ms, dom_n = gemm_kinds[dom]
        dom_flops = flops_per_gemm * gemms_in[dom]
        achieved = dom_flops / (dom_ms * 1e-3) / 1e12
        peak = peaks["sustained"]
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "dominant_kernel_traffic.json")
        # the ncu capture under profiles/ is of the default workload on one GPU: no figure for anything else
        profiled = (B, C, D, world) == (WORKLOAD["B"], WORKLOAD["C"], WORKLOAD["D"], 1)
        if profiled and os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get(dom)
            except Exception:
                traffic = None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "bf16", "data": "synthetic", "config": workload_config(args, world),
            "loss": final_loss,
            "step_tflops_algorithmic": 6.0 * B * C * D / (ms_per_step * 1e-3) / 1e12,
            "step_frac_of_peak": 6.0 * B * C * D / (ms_per_step * 1e-3) / 1e12 / (peak * world),
            "roofline": {
                "kernel": dom, "bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": f"{peaks['src']} (sustained bf16, MEASURED_PEAKS.json)",
                "launches_per_step": dom_n, "ms_per_step": dom_ms,
                "algorithmic_flops_per_step": dom_flops,
                "executed_flops_per_step": flops_per_gemm * {"bwd_persistent": 3, "bwd_g_gemm": 1}.get(dom, gemms_in[dom]),
            },
            "kernels_ms_per_step": {k: round(v[0], 4) for k, v in kinds.items()},
            "kernels_launches_per_step": {k: v[1] for k, v in kinds.items()},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": B * D * 4 + B * 8, "d2h_bytes_per_step": 4,
                    "ms_per_step": e2e_wall_ms / args.steps, "timer": "wall clock between device syncs"},
            "gpu_launches": int(launches),
            "clocks": clocks,
        }
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


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
