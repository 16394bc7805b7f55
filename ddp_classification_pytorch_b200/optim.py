"""Optimizer step for the ArcFace weight, fused with the next forward's ``F.normalize(self.weight)``.

The reference puts the head's parameter in the same ``torch.optim.SGD(lr, weight_decay=5e-4, momentum=0.9)`` /
``torch.optim.Adam(lr)`` as the backbone (ARCFACE/arc_main.py:248-253) and steps it once per batch (:330).  At a
million classes that step plus the re-normalisation at the top of the next forward (:159) is the largest HBM
consumer left once the head itself is fused.  :class:`HeadSGD` / :class:`HeadAdam` are ``torch.optim.Optimizer``
subclasses for the head's parameter only -- same hyper-parameters, same update rule, same ``state_dict`` layout
as the torch optimizers they replace (checkpoints interchange; ``lr_scheduler.StepLR`` (:255) works on them) --
that run ONE kernel per step: read w / dw / state, update, write w / state, and deposit the bf16 unit rows +
1/norm of the updated weight into the head's cached forward workspace, so the next ``forward_loss`` skips its
normalisation pass.

    backbone_opt = torch.optim.SGD(resnet50.parameters(), lr=LR, weight_decay=5e-4, momentum=0.9)
    head_opt = HeadSGD(ARC, lr=LR, weight_decay=5e-4, momentum=0.9)        # ARC: ArcMarginProduct (not the DDP wrapper)

All arithmetic is in ``libarcface_b200.so`` (``arcb_sgd_step`` / ``arcb_adam_step``); CUDA fp32 parameters only,
there is no CPU path.
"""
from __future__ import annotations

import torch
from torch.optim import Optimizer

from . import _lib
from .head import ArcMarginProduct


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


class _HeadOptimizer(Optimizer):
    def __init__(self, head: ArcMarginProduct, defaults):
        if not isinstance(head, ArcMarginProduct):
            raise TypeError("pass the ArcMarginProduct module itself (for a DDP-wrapped head: ARC.module)")
        super().__init__([head.weight], defaults)
        self._head = head
        head.cache_normalized_weight(True)

    def _prepare(self, p):
        """-> (grad, w_hat ptr | None, inv_norm ptr | None, Dp, cache) after validating the parameter."""
        if not p.is_cuda:
            raise RuntimeError("fused head optimizer: the weight must live on a CUDA device (no CPU path)")
        if p.dtype != torch.float32 or not p.is_contiguous():
            raise RuntimeError("fused head optimizer: the weight must be a contiguous fp32 tensor")
        g = p.grad
        if g.is_sparse:
            raise RuntimeError("fused head optimizer does not support sparse gradients")
        g = g.detach()
        if g.dtype != torch.float32 or not g.is_contiguous():
            g = g.float().contiguous()
        L = _lib.lib()
        rows, D = p.shape
        cache = self._head._wcache
        what = inv = None
        if (cache is not None and cache.busy == 0 and cache.ws is not None and cache.key == (rows, D, p.device)
                and p is self._head.weight):
            what, inv = cache.slots()
        if cache is not None:
            cache.valid = None  # the weight is about to change
        return g, what, inv, L.arcb_padded_dim(D), (cache if what is not None else None)

    def _commit(self, p, cache):
        torch.autograd.graph.increment_version(p)  # the kernel wrote through a raw pointer
        if cache is not None:
            cache.valid = (p.data_ptr(), p._version)


class HeadSGD(_HeadOptimizer):
    """``torch.optim.SGD`` for the head's weight (arc_main.py:252-253), one fused kernel per step."""

    def __init__(self, head, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False):
        if lr < 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if momentum < 0.0:
            raise ValueError(f"Invalid momentum value: {momentum}")
        if weight_decay < 0.0:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        if nesterov and (momentum <= 0 or dampening != 0):
            raise ValueError("Nesterov momentum requires a momentum and zero dampening")
        super().__init__(head, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                    nesterov=nesterov))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        L = _lib.lib()
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                g, what, inv, Dp, cache = self._prepare(p)
                rows, D = p.shape
                state = self.state[p]
                buf = state.get("momentum_buffer")
                first = buf is None
                if group["momentum"] != 0 and first:
                    buf = state["momentum_buffer"] = torch.empty_like(p, memory_format=torch.contiguous_format)
                _lib.check(L.arcb_sgd_step(p.data_ptr(), D, g.data_ptr(), D, None if buf is None else buf.data_ptr(), D,
                                           rows, D, float(group["lr"]), float(group["momentum"]),
                                           float(group["dampening"]), float(group["weight_decay"]),
                                           int(bool(group["nesterov"])), int(first), what, Dp, inv, _stream()),
                           "arcb_sgd_step")
                self._commit(p, cache)
        return loss


class HeadAdam(_HeadOptimizer):
    """``torch.optim.Adam`` for the head's weight (arc_main.py:249-250), one fused kernel per step."""

    def __init__(self, head, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0):
        if lr < 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if eps < 0.0:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameters: {betas}")
        if weight_decay < 0.0:
            raise ValueError(f"Invalid weight_decay value: {weight_decay}")
        super().__init__(head, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        L = _lib.lib()
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                g, what, inv, Dp, cache = self._prepare(p)
                rows, D = p.shape
                state = self.state[p]
                if len(state) == 0:  # same state layout as torch.optim.Adam
                    state["step"] = torch.tensor(0.0, dtype=torch.float32)
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                state["step"] += 1
                b1, b2 = group["betas"]
                _lib.check(L.arcb_adam_step(p.data_ptr(), D, g.data_ptr(), D, state["exp_avg"].data_ptr(),
                                            state["exp_avg_sq"].data_ptr(), D, rows, D, float(group["lr"]), float(b1),
                                            float(b2), float(group["eps"]), float(group["weight_decay"]),
                                            int(state["step"].item()), what, Dp, inv, _stream()), "arcb_adam_step")
                self._commit(p, cache)
        return loss
