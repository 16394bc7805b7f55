#!/usr/bin/env python
"""End-to-end ResNet-50 + ArcFace training step on synthetic images (BASELINE.json configs[4]).

Mirrors the wiring of the reference trainer (ARCFACE/arc_main.py:223-255 model / optimizer, :320-330 step):
torchvision resnet50 whose fc is Linear(2048,512)-ReLU-Linear(512,256)-LogSoftmax, the ArcFace head on its
256-d output, SGD(momentum 0.9, wd 5e-4) over both parameter groups, CrossEntropy on the head.  The backbone
and its DDP gradient all-reduce stay plain torch / cuDNN / NCCL; only the head is this repository's code.

  python examples/e2e_train.py --classes 10000 --batch 128 --steps 20
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 examples/e2e_train.py --head sharded

--head replicated : the reference's scheme -- every rank holds the full head under DDP (arc_main.py:243),
                    dW [C,D] is all-reduced by DDP.
--head sharded    : classes sharded over the ranks; features / labels are all-gathered so every rank's shard sees
                    the global batch; only [B,4] statistics cross NVLink in forward, dX returns by reduce_scatter.
--head torch      : the reference head in eager fp32 PyTorch (oracle.TorchPort's op sequence), for comparison.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F


class EagerArcHead(nn.Module):
    """The reference head's op sequence in eager fp32 PyTorch (arc_main.py:157-176) -- comparison arm only."""

    def __init__(self, d, c, s=30.0, m=0.5, easy_margin=True):
        super().__init__()
        import math
        self.s, self.easy = s, easy_margin
        self.cos_m, self.sin_m = math.cos(m), math.sin(m)
        self.th, self.mm = math.cos(math.pi - m), math.sin(math.pi - m) * m
        self.weight = nn.Parameter(torch.empty(c, d))
        nn.init.xavier_uniform_(self.weight)

    def forward_loss(self, x, y):
        cos = F.linear(F.normalize(x), F.normalize(self.weight))
        sin = (1.0 - cos.pow(2)).clamp(0, 1).sqrt()
        phi = cos * self.cos_m - sin * self.sin_m
        phi = torch.where(cos > 0, phi, cos) if self.easy else torch.where(cos > self.th, phi, cos - self.mm)
        hot = torch.zeros_like(cos).scatter_(1, y.view(-1, 1), 1)
        return F.cross_entropy((hot * phi + (1 - hot) * cos) * self.s, y)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--classes", type=int, default=10000)
    ap.add_argument("--batch", type=int, default=128, help="per-GPU batch")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--head", default="replicated", choices=["replicated", "sharded", "torch"])
    ap.add_argument("--image", type=int, default=224)
    ap.add_argument("--channels-last", action="store_true")
    ap.add_argument("--fused-head-optimizer", action="store_true",
                    help="step the head with HeadSGD (one pass that also prepares the next forward's normalised weight)")
    ap.add_argument("--out", default="", help="append the JSON result line to this file as well")
    args = ap.parse_args()

    from ddp_classification_pytorch_b200.trainer import ArcFaceTrainer, reference_backbone

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    torch.backends.cudnn.benchmark = True  # arc_main.py:51
    torch.manual_seed(999)  # arc_main.py:61

    model = reference_backbone(pretrained=False).to(dev)  # arc_main.py:223-231 (pretrained weights need the network)
    if args.channels_last:
        model = model.to(memory_format=torch.channels_last)
    group = dist.group.WORLD if world > 1 else None
    trainer = ArcFaceTrainer(
        model, 256, args.classes, head="sharded" if args.head == "sharded" else "replicated", process_group=group,
        s=30.0, m=0.5, easy_margin=True,  # arc_main.py:234
        lr=0.001, momentum=0.9, weight_decay=5e-4, fused_head_optimizer=args.fused_head_optimizer,
        head_module=EagerArcHead(256, args.classes) if args.head == "torch" else None, device=dev)

    g = torch.Generator(device=dev).manual_seed(rank)
    x = torch.randn(args.batch, 3, args.image, args.image, device=dev, generator=g)
    if args.channels_last:
        x = x.to(memory_format=torch.channels_last)
    y = torch.randint(0, args.classes, (args.batch,), device=dev, generator=g)

    for _ in range(args.warmup):
        trainer.step(x, y)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = trainer.step(x, y)
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        line = json.dumps({
            "metric": "ResNet-50 + ArcFace end-to-end training images/s (synthetic 224x224)",
            "value": args.batch * world * args.steps / (ms.item() * 1e-3), "unit": "images/s", "n_gpus": world,
            "ms_per_step": ms.item() / args.steps, "head": args.head, "classes": args.classes,
            "global_batch": args.batch * world, "loss": float(loss),
            "loss_is": "mean over the global batch" if args.head == "sharded" and world > 1 else "mean over rank 0's rows",
            "backbone": "torchvision resnet50 fp32 (cuDNN), untouched", "data": "synthetic"})
        print(line)
        if args.out:
            with open(args.out, "a") as fh:
                fh.write(line + "\n")
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
