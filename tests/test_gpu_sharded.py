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
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_nccl_two_ranks_equal_unsharded():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    with ctx.Manager() as mgr:
        out = mgr.dict()
        procs = [ctx.Process(target=_nccl_worker, args=(r, 2, port, out)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(300)
            assert p.exitcode == 0
        assert out.get(0) is True and out.get(1) is True, dict(out)
