"""The reference's training step (arc_main.py:320-330) with this head inside, on a tiny backbone: the loss trajectory of
a few SGD steps must follow the same step taken with the reference head's eager fp32 op sequence."""
import math
import os
import socket

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

from ddp_classification_pytorch_b200.trainer import ArcFaceTrainer

pytestmark = pytest.mark.gpu


class EagerHead(nn.Module):
    """ArcMarginProduct.forward (arc_main.py:157-176) + CrossEntropyLoss in eager fp32 torch: the comparison arm."""

    def __init__(self, d, c, s=30.0, m=0.5, easy=True):
        super().__init__()
        self.s, self.m, self.easy = s, m, easy
        self.weight = nn.Parameter(torch.empty(c, d))

    def forward_loss(self, x, y):
        cos = F.linear(F.normalize(x), F.normalize(self.weight))
        sine = torch.sqrt((1.0 - torch.pow(cos, 2)).clamp(0, 1))
        phi = cos * math.cos(self.m) - sine * math.sin(self.m)
        phi = torch.where(cos > 0, phi, cos) if self.easy else torch.where(cos > math.cos(math.pi - self.m), phi,
                                                                            cos - math.sin(math.pi - self.m) * self.m)
        hot = torch.zeros_like(cos).scatter_(1, y.view(-1, 1), 1)
        return F.cross_entropy((hot * phi + (1 - hot) * cos) * self.s, y)


def _tiny_backbone(feat):
    torch.manual_seed(11)
    return nn.Sequential(nn.Conv2d(3, 8, 3, stride=2, padding=1), nn.BatchNorm2d(8), nn.ReLU(), nn.AdaptiveAvgPool2d(2),
                         nn.Flatten(), nn.Linear(32, feat), nn.LogSoftmax(dim=1))  # ends like the reference's fc (:226-230)


def _data(b, C, seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(b, 3, 16, 16, generator=g), torch.randint(0, C, (b,), generator=g)


def _head_weight(C, feat):
    g = torch.Generator().manual_seed(3)
    return (torch.rand(C, feat, generator=g) * 2 - 1) * 0.1


@pytest.mark.parametrize("fused_opt", [False, True])
def test_three_steps_follow_the_eager_head(fused_opt):
    feat, C, b = 32, 700, 48
    dev = torch.device("cuda", 0)
    w0 = _head_weight(C, feat)
    x, y = _data(b, C, 1)
    x, y = x.to(dev), y.to(dev)

    def run(eager):
        bb = _tiny_backbone(feat).to(dev)
        hm = EagerHead(feat, C) if eager else None
        tr = ArcFaceTrainer(bb, feat, C, head="replicated", lr=0.05, momentum=0.9, weight_decay=5e-4,
                            fused_head_optimizer=fused_opt and not eager, head_module=hm, device=dev)
        with torch.no_grad():
            tr.head_core.weight.copy_(w0)
        return [tr.step(x, y).item() for _ in range(4)], tr.head_core.weight.detach().clone()

    ref, w_ref = run(True)
    got, w_got = run(False)
    assert ref[-1] < ref[0]  # it trains
    assert got == pytest.approx(ref, rel=2e-3)
    assert torch.nn.functional.cosine_similarity((w_got - w0.to(dev)).flatten(), (w_ref - w0.to(dev)).flatten(), dim=0) > 0.995


def _ddp_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import torch.distributed as dist
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        feat, C, b = 32, 900, 32
        w0 = _head_weight(C, feat)
        x, y = _data(b, C, 100 + rank)  # every rank trains on its own rows
        x, y = x.to(dev), y.to(dev)
        traj = {}
        for mode in ("replicated", "sharded"):
            tr = ArcFaceTrainer(_tiny_backbone(feat).to(dev), feat, C, head=mode, process_group=dist.group.WORLD, lr=0.05,
                                device=dev)
            with torch.no_grad():
                lo, hi = (tr.head_core.class_lo, tr.head_core.class_hi) if mode == "sharded" else (0, C)
                tr.head_core.weight.copy_(w0[lo:hi])
            losses = []
            for _ in range(3):
                l = tr.step(x, y)
                if mode == "replicated":  # per-rank mean -> the global mean the sharded head reports
                    dist.all_reduce(l)
                    l = l / world
                losses.append(l.item())
            w = tr.head_core.full_weight() if mode == "sharded" else tr.head_core.weight.detach().clone()
            p0 = next(tr.backbone.parameters()).detach().clone()
            traj[mode] = (losses, w, p0)
        (l_r, w_r, p_r), (l_s, w_s, p_s) = traj["replicated"], traj["sharded"]
        ok = (l_s == pytest.approx(l_r, rel=2e-3)
              and torch.allclose(w_s, w_r, rtol=1e-2, atol=2e-5)          # same head updates, with no dW all-reduce
              and torch.allclose(p_s, p_r, rtol=1e-2, atol=2e-5))         # same backbone updates (dX x world, DDP mean)
        out[rank] = bool(ok)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_and_replicated_trainers_take_the_same_steps(world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    with ctx.Manager() as mgr:
        out = mgr.dict()
        procs = [ctx.Process(target=_ddp_worker, args=(r, world, port, out)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(300)
            assert p.exitcode == 0
        assert all(out.get(r) is True for r in range(world)), dict(out)
