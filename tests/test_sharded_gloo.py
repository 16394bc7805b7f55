"""Class-sharded path on CPU: world_size-2 (and 3) gloo process groups.

The module's collective plumbing (shard bounds, class offsets, all_gather of the [B,4] shard
records, merge, all_reduce of the dX_hat partials, full_weight gather) is exercised for real; the
per-shard arithmetic -- which in the product is CUDA only -- is supplied by a checker engine built
on the oracle.  The result must equal the un-sharded oracle.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ddp_classification_pytorch_b200 import ArcMarginProduct, shard_bounds
from ddp_classification_pytorch_b200.head import _canonical_inputs
from oracle import arcface_oracle as orc


class OracleEngine:
    """Engine with the interface of head.CudaEngine, computing each shard with the CPU oracle."""

    def prepare(self, x, w, label):
        return _canonical_inputs(x, w, label)

    def shard_forward(self, xc, wc, lab, cfg):
        mx, se, tgt = orc.shard_forward_stats(xc.numpy(), wc.numpy(), lab.numpy(), cfg["class_offset"], cfg["s"],
                                              cfg["m"], cfg["easy_margin"])
        stats = torch.tensor(np.stack([mx, se, tgt, np.zeros_like(mx)], axis=1))  # float64 record
        return stats, (xc, wc, lab)

    def shard_top3(self, xc, wc, state, cfg):
        # margin only where this shard owns the label
        lo = cfg["class_offset"]
        lab = state[2].numpy() - lo
        cos_logits = orc.head_forward(xc.numpy(), wc.numpy(), np.clip(lab, 0, wc.shape[0] - 1), cfg["s"], cfg["m"],
                                      cfg["easy_margin"])
        plain = cfg["s"] * (orc.l2_normalize_rows(xc.numpy())[0] @ orc.l2_normalize_rows(wc.numpy())[0].T)
        owned = (lab >= 0) & (lab < wc.shape[0])
        z = np.where(owned[:, None], cos_logits, plain)
        order = np.argsort(-z, axis=1, kind="stable")[:, :3]
        val = np.take_along_axis(z, order, axis=1)
        if order.shape[1] < 3:  # shard with fewer than three classes
            pad = 3 - order.shape[1]
            order = np.pad(order, ((0, 0), (0, pad)), constant_values=-1 - lo)
            val = np.pad(val, ((0, 0), (0, pad)), constant_values=-np.inf)
        return torch.tensor(val), torch.tensor(order + lo)

    def merge_loss(self, gathered, P, B):
        g = gathered.reshape(P, B, 4).numpy()
        lse, tgt, loss = orc.merge_shard_stats(g[:, :, 0], g[:, :, 1], g[:, :, 2])
        return torch.tensor(lse), torch.tensor(tgt), torch.tensor(loss, dtype=torch.float32)

    def shard_backward(self, state, cfg, lse=None, zt=None, grad_loss=None, grad_logits=None):
        xc, wc, lab = state
        dw, dxh = orc.shard_backward(xc.numpy(), wc.numpy(), lab.numpy(), cfg["class_offset"], lse.numpy(), cfg["s"],
                                     cfg["m"], cfg["easy_margin"], float(grad_loss))
        return torch.tensor(dw), torch.tensor(dxh)

    def input_grad(self, xc, dxhat):
        return torch.tensor(orc.input_grad(xc.numpy(), dxhat.numpy()))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, B, D, C, easy, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(123)
        w_full = (torch.rand(C, D, generator=g) * 2 - 1) * 0.1
        x = torch.randn(B, D, generator=g)
        y = torch.randint(0, C, (B,), generator=g)
        head = ArcMarginProduct(D, C, s=30.0, m=0.5, easy_margin=easy, process_group=dist.group.WORLD,
                                _engine=OracleEngine())
        lo, hi = shard_bounds(C, world, rank)
        assert (head.class_lo, head.class_hi) == (lo, hi) and tuple(head.weight.shape) == (hi - lo, D)
        with torch.no_grad():
            head.weight.copy_(w_full[lo:hi])
        xg = x.clone().requires_grad_(True)
        loss = head.forward_loss(xg, y, want_top3=True)
        (loss * 2.0).backward()
        full = head.full_weight()
        ref = orc.head_forward_backward(x.numpy(), w_full.numpy(), y.numpy(), 30.0, 0.5, easy, grad_scale=2.0)
        ok = (
            abs(loss.item() - ref["loss"]) < 1e-5 * abs(ref["loss"])
            and np.allclose(xg.grad.numpy(), ref["dx"], rtol=1e-4, atol=1e-7)
            and np.allclose(head.weight.grad.numpy(), ref["dw"][lo:hi], rtol=1e-4, atol=1e-7)
            and torch.equal(full, w_full)
            and np.allclose(head.aux["lse"].numpy(), ref["lse"], rtol=1e-6)
            and np.array_equal(head.aux["top3"].numpy(), ref["top3"])
            and np.array_equal(head.aux["argmax"].numpy(), ref["argmax"])
        )
        # accuracy counters (arc_main.py:338-345): sharded mode scores the same global batch on every rank
        counts = head.correct_counts(y)
        top3 = ref["top3"]
        ok = ok and counts.tolist() == [int((top3[:, 0] == y.numpy()).sum()),
                                        int((top3 == y.numpy()[:, None]).any(1).sum()), B]
        # data-parallel counters: each rank scored its own samples -> summed over the group
        summed = head.correct_counts(y, reduce_group=dist.group.WORLD)
        ok = ok and summed.tolist() == [world * c for c in counts.tolist()]
        # checkpoint interchange with the reference layout: full weight out, full weight in (other row order)
        sd = head.full_state_dict()
        ok = ok and list(sd) == ["weight"] and torch.equal(sd["weight"], w_full)
        head.load_full_state_dict({"weight": torch.flip(w_full, dims=[0])})
        ok = ok and torch.equal(head.weight.detach(), torch.flip(w_full, dims=[0])[lo:hi])
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,easy", [(2, False), (2, True), (3, False)])
def test_sharded_head_equals_unsharded_oracle(world, easy):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        out = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, world, port, 24, 32, 101, easy, out)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        assert all(out.get(r) is True for r in range(world)), dict(out)


# ---- the reference's own multi-GPU wiring: DistributedDataParallel(ARC), arc_main.py:243 -------------------------
def _ddp_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        B, D, C = 12, 16, 37
        g = torch.Generator().manual_seed(77)
        w_full = (torch.rand(C, D, generator=g) * 2 - 1) * 0.1
        xs = [torch.randn(B, D, generator=g) for _ in range(world)]      # every rank scores its OWN batch
        ys = [torch.randint(0, C, (B,), generator=g) for _ in range(world)]
        head = ArcMarginProduct(D, C, s=30.0, m=0.5, easy_margin=True, _engine=OracleEngine())
        with torch.no_grad():
            head.weight.copy_(w_full)
        arc = torch.nn.parallel.DistributedDataParallel(head)            # arc_main.py:243
        xg = xs[rank].clone().requires_grad_(True)
        loss = arc(xg, ys[rank], fused_loss=True)                        # through the wrapper: the reducer is armed
        loss.backward()                                                  # arc_main.py:329
        refs = [orc.head_forward_backward(xs[r].numpy(), w_full.numpy(), ys[r].numpy(), 30.0, 0.5, True) for r in range(world)]
        mean_dw = sum(r["dw"] for r in refs) / world                     # what DDP's all-reduce (average) must leave
        grads = [torch.empty_like(head.weight.grad) for _ in range(world)]
        dist.all_gather(grads, head.weight.grad.contiguous())
        ok = (
            abs(loss.item() - refs[rank]["loss"]) < 1e-5 * abs(refs[rank]["loss"])
            and all(torch.equal(grads[0], gr) for gr in grads)           # replicas hold the SAME gradient
            and np.allclose(head.weight.grad.numpy(), mean_dw, rtol=1e-4, atol=1e-8)
            and np.allclose(xg.grad.numpy(), refs[rank]["dx"], rtol=1e-4, atol=1e-8)
        )
        # the trap this entry exists for: the module's own method bypasses the wrapper, nothing is reduced
        head.weight.grad = None
        head.forward_loss(xs[rank], ys[rank]).backward()
        dist.all_gather(grads, head.weight.grad.contiguous())
        ok = ok and not torch.equal(grads[0], grads[1])
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


def test_fused_loss_through_ddp_wrapper_reduces_weight_grad():
    world = 2
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        out = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_ddp_worker, args=(r, world, port, out)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        assert all(out.get(r) is True for r in range(world)), dict(out)


# ---- class-sharded head behind a data-parallel backbone: all_gather of the batch, reduce_scatter of dX -------------
def _gather_worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        b, D, C = 10, 16, 53
        g = torch.Generator().manual_seed(5)
        w_full = (torch.rand(C, D, generator=g) * 2 - 1) * 0.1
        x_all = torch.randn(world * b, D, generator=g)
        y_all = torch.randint(0, C, (world * b,), generator=g)
        head = ArcMarginProduct(D, C, s=30.0, m=0.5, easy_margin=False, process_group=dist.group.WORLD,
                                _engine=OracleEngine())
        lo, hi = head.class_lo, head.class_hi
        with torch.no_grad():
            head.weight.copy_(w_full[lo:hi])
        rows = slice(rank * b, (rank + 1) * b)
        xg = x_all[rows].clone().requires_grad_(True)
        loss = head.forward_loss(xg, y_all[rows], want_top3=True, gather_batch=True)
        loss.backward()
        ref = orc.head_forward_backward(x_all.numpy(), w_full.numpy(), y_all.numpy(), 30.0, 0.5, False)
        counts = head.correct_counts()  # scores the gathered labels kept in aux
        top3 = ref["top3"]
        ok = (
            abs(loss.item() - ref["loss"]) < 1e-5 * abs(ref["loss"])                      # mean over the GLOBAL batch
            and tuple(xg.grad.shape) == (b, D)
            and np.allclose(xg.grad.numpy(), world * ref["dx"][rows], rtol=1e-4, atol=1e-8)  # own rows, x world (DDP averages)
            and np.allclose(head.weight.grad.numpy(), ref["dw"][lo:hi], rtol=1e-4, atol=1e-8)
            and torch.equal(head.aux["labels"], y_all)
            and counts.tolist() == [int((top3[:, 0] == y_all.numpy()).sum()),
                                    int((top3 == y_all.numpy()[:, None]).any(1).sum()), world * b]
        )
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_gather_batch_reduce_scatter_equals_global_oracle(world):
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        out = mgr.dict()
        port = _free_port()
        procs = [ctx.Process(target=_gather_worker, args=(r, world, port, out)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        assert all(out.get(r) is True for r in range(world)), dict(out)
