"""Class-sharded CUDA path.

* one GPU: the shards are run one after the other through the C ABI (class offsets, "label not in
  this shard" records, P-way merge, summed dX_hat partials) and compared with the un-sharded oracle;
* two or more GPUs: the real thing, one process per GPU over NCCL (skipped on a single-GPU box).
"""
import os
import socket

import numpy as np
import pytest
import torch

from conftest import cosine_similarity
from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib, shard_bounds
from ddp_classification_pytorch_b200.head import CudaEngine
from oracle import arcface_oracle as orc

pytestmark = pytest.mark.gpu


def _problem(B, D, C, seed=5):
    g = torch.Generator().manual_seed(seed)
    w = (torch.rand(C, D, generator=g) * 2 - 1) * 0.08
    x = torch.randn(B, D, generator=g)
    y = torch.randint(0, C, (B,), generator=g)
    return x, w, y


@pytest.mark.parametrize("parts", [2, 3, 8])
@pytest.mark.parametrize("easy", [False, True])
def test_sequential_shards_equal_unsharded(parts, easy):
    B, D, C = 160, 128, 2500
    x, w, y = _problem(B, D, C)
    ref = orc.head_forward_backward(x.numpy(), w.numpy(), y.numpy(), 30.0, 0.5, easy)
    eng = CudaEngine()
    xc, yc = x.cuda(), y.cuda()
    recs, states, spans = [], [], []
    for r in range(parts):
        lo, hi = shard_bounds(C, parts, r)
        cfg = {"s": 30.0, "m": 0.5, "easy_margin": easy, "class_offset": lo, "chunk_classes": 512}
        stats, state = eng.shard_forward(xc, w[lo:hi].cuda().contiguous(), yc, cfg)
        recs.append(stats); states.append((state, cfg)); spans.append((lo, hi))
    gathered = torch.stack(recs).contiguous()
    # exactly one shard owns each label
    owners = torch.isfinite(gathered[:, :, 2]).sum(0)
    assert (owners == 1).all()
    lse, zt, loss = eng.merge_loss(gathered, parts, B)
    assert loss.item() == pytest.approx(ref["loss"], rel=1e-3)
    np.testing.assert_allclose(lse.cpu().numpy(), ref["lse"], rtol=2e-3)
    np.testing.assert_allclose(zt.cpu().numpy(), ref["target_logit"], atol=1e-3, rtol=1e-5)
    one = torch.ones((), device="cuda")
    dxhat_sum, dws = None, []
    for (state, cfg) in states:
        dw, dxhat = eng.shard_backward(state, cfg, lse=lse, zt=zt, grad_loss=one)
        dws.append(dw)
        dxhat_sum = dxhat if dxhat_sum is None else dxhat_sum + dxhat
    dx = eng.input_grad(xc, dxhat_sum)
    _lib.check(_lib.lib().arcb_check_watchdog(torch.cuda.current_stream().cuda_stream))
    assert cosine_similarity(dx.cpu().numpy(), ref["dx"]) >= 0.999
    assert cosine_similarity(torch.cat(dws).cpu().numpy(), ref["dw"]) >= 0.999


def _nccl_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        B, D, C = 256, 256, 6000
        x, w, y = _problem(B, D, C, seed=9)
        head = ArcMarginProduct(D, C, s=30.0, m=0.5, easy_margin=True, process_group=dist.group.WORLD).cuda()
        lo, hi = head.class_lo, head.class_hi
        with torch.no_grad():
            head.weight.copy_(w[lo:hi])
        xg = x.cuda().requires_grad_(True)
        loss = head.forward_loss(xg, y.cuda(), want_top3=True)
        loss.backward()
        torch.cuda.synchronize()
        ref = orc.head_forward_backward(x.numpy(), w.numpy(), y.numpy(), 30.0, 0.5, True)
        top = np.take_along_axis(ref["logits"], ref["top3"], axis=1)
        clear = ((top[:, :-1] - top[:, 1:]) > 1e-3).all(axis=1)
        ok = (abs(loss.item() - ref["loss"]) <= 1e-3 * abs(ref["loss"])
              and cosine_similarity(xg.grad.cpu().numpy(), ref["dx"]) >= 0.999
              and cosine_similarity(head.weight.grad.cpu().numpy(), ref["dw"][lo:hi]) >= 0.999
              and torch.equal(head.full_weight().cpu(), w)
              and clear.mean() > 0.9
              and np.array_equal(head.aux["top3"].cpu().numpy()[clear], ref["top3"][clear]))
        # the statistics exchange done by the kernels over symmetric (NVLink peer) memory must have been used, and
        # must give exactly what the NCCL all_gather path gives (same records, same merge arithmetic)
        peer_used = bool(head._peer) and not head._peer_off
        nccl_head = ArcMarginProduct(D, C, s=30.0, m=0.5, easy_margin=True, process_group=dist.group.WORLD,
                                     peer_exchange=False).cuda()
        with torch.no_grad():
            nccl_head.weight.copy_(w[lo:hi])
        for _ in range(3):  # several steps: both slots of the double-buffered gather buffer, growing epochs
            xn = x.cuda().requires_grad_(True)
            ln = nccl_head.forward_loss(xn, y.cuda())
            xp = x.cuda().requires_grad_(True)
            lp = head.forward_loss(xp, y.cuda())
            ok = ok and abs(lp.item() - ln.item()) <= 1e-6 * abs(ln.item()) and torch.equal(head.aux["lse"], nccl_head.aux["lse"])
            # backward: partial sums of dX_hat reduced by the kernel over peer memory vs NCCL all_reduce + row kernel
            head.weight.grad = None
            nccl_head.weight.grad = None
            lp.backward()
            ln.backward()
            torch.cuda.synchronize()
            ok = (ok and torch.allclose(xp.grad, xn.grad, rtol=1e-4, atol=1e-9)       # other summation order over ranks
                  and torch.allclose(head.weight.grad, nccl_head.weight.grad, rtol=1e-3, atol=1e-7))
        out[f"peer{rank}"] = peer_used
        head.weight.grad = None
        # data-parallel rows in, reduce_scatter of dX out: this rank's rows only
        b = B // world
        rows = slice(rank * b, (rank + 1) * b)
        head.weight.grad = None
        xl = x[rows].cuda().requires_grad_(True)
        loss2 = head.forward_loss(xl, y[rows].cuda(), gather_batch=True)
        loss2.backward()
        torch.cuda.synchronize()
        n = world * b
        ref2 = orc.head_forward_backward(x[:n].numpy(), w.numpy(), y[:n].numpy(), 30.0, 0.5, True)
        ok = (ok and abs(loss2.item() - ref2["loss"]) <= 1e-3 * abs(ref2["loss"])
              and tuple(xl.grad.shape) == (b, D)
              and cosine_similarity(xl.grad.cpu().numpy(), ref2["dx"][rows]) >= 0.999
              and np.linalg.norm(xl.grad.cpu().numpy()) == pytest.approx(world * np.linalg.norm(ref2["dx"][rows]), rel=2e-2)
              and cosine_similarity(head.weight.grad.cpu().numpy(), ref2["dw"][lo:hi]) >= 0.999)
        # the reference's wiring: a replicated head under DDP, fused loss through the wrapper -> dW all-reduced
        rep = ArcMarginProduct(D, C, s=30.0, m=0.5, easy_margin=True).cuda()
        with torch.no_grad():
            rep.weight.copy_(w)
        arc = torch.nn.parallel.DistributedDataParallel(rep, device_ids=[rank])
        arc(x[rows].cuda(), y[rows].cuda(), fused_loss=True).backward()
        torch.cuda.synchronize()
        per_rank = [orc.head_forward_backward(x[r * b:(r + 1) * b].numpy(), w.numpy(), y[r * b:(r + 1) * b].numpy(),
                                              30.0, 0.5, True)["dw"] for r in range(world)]
        ok = ok and cosine_similarity(rep.weight.grad.cpu().numpy(), sum(per_rank) / world) >= 0.999
        grads = [torch.empty_like(rep.weight.grad) for _ in range(world)]
        dist.all_gather(grads, rep.weight.grad.contiguous())
        ok = ok and all(torch.equal(grads[0], gr) for gr in grads)
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4, 8])
def test_nccl_ranks_equal_unsharded(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    with ctx.Manager() as mgr:
        out = mgr.dict()
        procs = [ctx.Process(target=_nccl_worker, args=(r, world, port, out)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(300)
            assert p.exitcode == 0
        assert all(out.get(r) is True for r in range(world)), dict(out)
        # symmetric memory is expected on an NVLink box; if torch could not set it up the NCCL path was tested instead
        if not all(out.get(f"peer{r}") for r in range(world)):
            import warnings
            warnings.warn("peer-memory statistics exchange was not available: NCCL all_gather path used")
