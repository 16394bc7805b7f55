"""The reference's training step around the ArcFace head (BASELINE.json configs[4]).

``ArcFaceTrainer`` wires a backbone, this package's ``ArcMarginProduct`` and the optimizer the way
ARCFACE/arc_main.py does (:223-255 model / optimizer construction, :320-330 one step):

    features = model(inputs)                      # :325   backbone, DDP + SyncBatchNorm when world > 1 (:239-240)
    loss = CE(ARC(features, labels), labels)      # :326-327  -- here ONE fused call, the logits are never written
    loss.backward(); optimizer.step()             # :329-330

The backbone, its DDP gradient all-reduce, SyncBatchNorm and the SGD step stay plain torch / cuDNN / NCCL (the
north_star leaves them untouched); only the head is this repository's code.  Two ways to place the head on
several GPUs:

``head="replicated"``  the reference's scheme (arc_main.py:243): every rank holds the full ``[C, D]`` weight under
    ``DistributedDataParallel``; the fused loss is reached THROUGH the wrapper (``ARC(f, y, fused_loss=True)``) so
    DDP all-reduces ``dW``.  Each rank's loss is the mean over its own rows, as in the reference.
``head="sharded"``     classes sharded over the ranks (``process_group``): ``forward_loss(gather_batch=True)``
    all-gathers features / labels, only ``[B, 4]`` statistics cross NVLink in forward, and dX comes back by
    ``reduce_scatter``.  The loss is the mean over the GLOBAL batch (identical on every rank); ``dW`` needs no
    collective at all.  Both schemes produce the same parameter updates.

There is no CPU path: the head's kernels need a CUDA device.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .head import ArcMarginProduct
from .optim import HeadSGD


def reference_backbone(pretrained: bool = False) -> nn.Module:
    """torchvision ResNet-50 with the reference's replacement ``fc`` (arc_main.py:223-231):
    Linear(2048, 512) - ReLU - Linear(512, 256) - LogSoftmax.  ``pretrained`` needs the network (arc_main.py:223
    uses it); synthetic benchmarks run from random initialisation."""
    import torchvision
    model = torchvision.models.resnet50(weights="IMAGENET1K_V1" if pretrained else None)
    model.fc = nn.Sequential(nn.Linear(model.fc.in_features, 512), nn.ReLU(), nn.Linear(512, 256), nn.LogSoftmax(dim=1))
    return model


class ArcFaceTrainer:
    """One training step of backbone + ArcFace head + SGD, replicated or class-sharded head.

    Args:
        backbone: module mapping a batch to ``[b, feat_dim]`` features (already on ``device``)
        feat_dim, num_classes: the head's ``in_features`` / ``out_features`` (arc_main.py:234: 256, 2173)
        head: ``"replicated"`` | ``"sharded"`` (see the module docstring)
        process_group: the ranks of the job (``None`` = single process)
        s, m, easy_margin: head hyper-parameters (arc_main.py:234: 30, 0.5, True)
        lr, momentum, weight_decay: ``torch.optim.SGD`` over both parameter groups (arc_main.py:252-253)
        fused_head_optimizer: step the head with ``HeadSGD`` (one kernel that also prepares the next forward's
            normalised weight) instead of putting its parameter into the torch optimizer
        head_module: a ready-made head to use instead (comparison arms in tests / examples); must offer
            ``forward_loss(features, labels)`` and ``parameters()``
    """

    def __init__(self, backbone: nn.Module, feat_dim: int, num_classes: int, *, head: str = "replicated",
                 process_group=None, s: float = 30.0, m: float = 0.5, easy_margin: bool = True, lr: float = 1e-3,
                 momentum: float = 0.9, weight_decay: float = 5e-4, fused_head_optimizer: bool = False,
                 head_module: Optional[nn.Module] = None, device=None):
        if head not in ("replicated", "sharded"):
            raise ValueError("head must be 'replicated' or 'sharded'")
        import torch.distributed as dist
        self.group = process_group
        self.world = dist.get_world_size(process_group) if process_group is not None else 1
        self.mode = head
        dev = device if device is not None else next(backbone.parameters()).device
        if torch.device(dev).type != "cuda":
            raise RuntimeError("ArcFaceTrainer needs a CUDA device (the head has no CPU path)")
        self.device = torch.device(dev)
        sharded = head == "sharded" and self.world > 1
        if head_module is not None:
            self.head_core = head_module.to(self.device)
        else:
            self.head_core = ArcMarginProduct(feat_dim, num_classes, s=s, m=m, easy_margin=easy_margin,
                                              process_group=process_group if sharded else None).to(self.device)
        self.sharded = sharded and head_module is None
        self.backbone = backbone
        self.head = self.head_core
        if self.world > 1:
            self.backbone = nn.SyncBatchNorm.convert_sync_batchnorm(self.backbone)  # arc_main.py:239
            self.backbone = nn.parallel.DistributedDataParallel(self.backbone, device_ids=[self.device.index],
                                                                process_group=process_group)  # :240
            if not self.sharded:  # arc_main.py:243 -- a sharded head owns its rows: nothing to all-reduce
                self.head = nn.parallel.DistributedDataParallel(self.head_core, device_ids=[self.device.index],
                                                                process_group=process_group)
        self.head_opt = None
        if fused_head_optimizer and isinstance(self.head_core, ArcMarginProduct):
            self.opt = torch.optim.SGD(self.backbone.parameters(), lr=lr, momentum=momentum, weight_decay=weight_decay)
            self.head_opt = HeadSGD(self.head_core, lr=lr, momentum=momentum, weight_decay=weight_decay)
        else:
            self.opt = torch.optim.SGD([{"params": self.backbone.parameters()}, {"params": self.head_core.parameters()}],
                                       lr=lr, momentum=momentum, weight_decay=weight_decay)  # arc_main.py:252-253

    def loss(self, images, labels, want_top3: bool = False):
        feats = self.backbone(images)  # arc_main.py:325
        if not isinstance(self.head_core, ArcMarginProduct):
            return self.head_core.forward_loss(feats, labels)
        if self.sharded:
            return self.head_core.forward_loss(feats, labels, want_top3=want_top3, gather_batch=True)
        # through the (possibly DDP-wrapped) module's __call__, so that DDP's reducer sees the backward
        return self.head(feats, labels, fused_loss=True, want_top3=want_top3)  # arc_main.py:326-327 fused

    def step(self, images, labels, want_top3: bool = False):
        """zero_grad / forward / backward / optimizer step (arc_main.py:324-330); returns the detached loss."""
        self.opt.zero_grad(set_to_none=True)
        if self.head_opt is not None:
            self.head_opt.zero_grad(set_to_none=True)
        loss = self.loss(images, labels, want_top3=want_top3)
        loss.backward()  # arc_main.py:329
        self.opt.step()  # arc_main.py:330
        if self.head_opt is not None:
            self.head_opt.step()
        return loss.detach()
