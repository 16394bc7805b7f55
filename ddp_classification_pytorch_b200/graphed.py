"""One CUDA graph for the head's whole training step (forward + loss + backward) at fixed shapes.

The reference loop (ARCFACE/arc_main.py:320-330) calls the head with the same batch shape every iteration.  For a
head whose step is short (up to ~100k classes per GPU: <= 0.6 ms of GPU work spread over eight kernels) launch gaps
and Python overhead are a visible share of it; replaying a captured graph removes them (measured on B200, B=1024,
D=512, 125k classes: 0.558 -> 0.506 ms per step).  For long steps it does not pay (1M classes: 4.23 eager vs
4.36 ms replayed), so this is an opt-in helper, not the default path.  Nothing is cached: every replay runs every
kernel -- including ``F.normalize(self.weight)`` -- on whatever the weight and the static input buffers hold at that
moment.  (A head with ``cache_normalized_weight`` / a fused optimizer keeps its cache for eager calls, but the
captured step is recorded with the cache switched off: a replay cannot re-decide "is the cached copy still valid?",
and a frozen "yes" would silently train on a stale normalised weight with any optimizer that does not refresh it.)  Single-process heads only (replicated mode): a
class-sharded head would put its two NCCL collectives inside the graph, which is not supported here.

    step = GraphedHeadStep(ARC, batch=1024)        # ARC: ArcMarginProduct on a CUDA device
    for features, labels in loader:
        loss = step(features, labels)              # copies into the static buffers, replays the graph
        # ARC.weight.grad and step.x.grad are static tensors that the replay has just filled
        optimizer.step()

Before constructing it, drop every reference to losses of earlier eager steps (an autograd graph kept alive from
another stream pins the weight's gradient accumulation to that stream and invalidates the capture).
Limits: fixed batch size; ``features`` arrives detached (its gradient is ``step.x.grad``: feed it to
``features.backward(step.x.grad)`` when a backbone sits in front); the module's hyper-parameters are frozen at
capture time.  Falls back to nothing: if the capture fails the constructor raises.
"""
from __future__ import annotations

import torch

from .head import ArcMarginProduct


class GraphedHeadStep:
    def __init__(self, head: ArcMarginProduct, batch: int, *, want_top3: bool = False, warmup: int = 3):
        if not isinstance(head, ArcMarginProduct):
            raise TypeError("GraphedHeadStep needs the ArcMarginProduct module itself")
        dev = head.weight.device
        if dev.type != "cuda":
            raise RuntimeError("GraphedHeadStep needs the head on a CUDA device (there is no CPU path)")
        if head.process_group is not None and head.world_size > 1:
            raise RuntimeError("GraphedHeadStep does not support a class-sharded head (collectives inside the graph)")
        self.head = head
        self.want_top3 = bool(want_top3)
        self.x = torch.zeros(batch, head.in_features, device=dev, dtype=torch.float32, requires_grad=True)
        self.y = torch.zeros(batch, device=dev, dtype=torch.int64)
        # eager warm-up on a side stream (lazy one-time set-up inside the library, NCCL communicators, allocator pools)
        saved_cache, head._wcache = head._wcache, None  # the captured forward always normalises the weight itself
        try:
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                for _ in range(max(1, warmup)):
                    head.weight.grad = None
                    self.x.grad = None
                    head.forward_loss(self.x, self.y, want_top3=self.want_top3).backward()
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            head.weight.grad = None
            self.x.grad = None
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.loss = head.forward_loss(self.x, self.y, want_top3=self.want_top3)
                self.loss.backward()
        finally:
            head._wcache = saved_cache
        self.aux = dict(head.aux)  # static tensors as well (lse, target_logit, top3 ...)

    def __call__(self, features=None, labels=None):
        """Copy ``features`` / ``labels`` (device or pinned host tensors) into the static buffers and replay."""
        if features is not None:
            self.x.data.copy_(features.detach(), non_blocking=True)
        if labels is not None:
            self.y.copy_(labels.detach().reshape(-1), non_blocking=True)
        self.graph.replay()
        return self.loss
