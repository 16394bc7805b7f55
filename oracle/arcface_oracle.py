"""CPU oracle for the ArcFace head hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this module.  The product package
(``ddp_classification_pytorch_b200``) never imports it and has no CPU fallback.

What it restates (all citations relative to /root/reference):

* ``ArcMarginProduct.__init__``   ARCFACE/arc_main.py:139-152  (constants cos_m, sin_m, th, mm)
* ``ArcMarginProduct.forward``    ARCFACE/arc_main.py:157-176  (normalise, cosine, margin, one-hot blend, scale)
* ``nn.CrossEntropyLoss()``       ARCFACE/arc_main.py:245, applied at :327 / :369 (mean reduction)
* ``loss.backward()``             ARCFACE/arc_main.py:329      (closed form of what autograd derives)
* logits consumers                ARCFACE/arc_main.py:338-345, :377-388 (argmax / top-3)

The arithmetic itself lives in third-party PyTorch (ATen), which the reference does not
pin (requirements.txt lists no torch); the effective pin is the torch in this image
(2.11.0+cu128).  Parity pin: the reference ships no tests / golden vectors, so this
restatement is pinned against the *reference class itself*, AST-extracted from
/root/reference and executed in the build container (see ``oracle/ref_extract.py`` and
``tests/golden/make_golden.py``); the resulting vectors are committed under
``tests/golden/``.

Two restatements are provided:

* :func:`head_forward_backward` -- numpy float64 closed form (the checker used by parity tests).
* :class:`TorchPort` -- a float32 torch-CPU port that performs the same sequence of dense
  B x C passes as the reference and differentiates with autograd; used only as the timed
  CPU baseline (``cpu_baseline.kind == "port"``).
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

NORMALIZE_EPS = 1e-12  # F.normalize default eps (arc_main.py:159 uses the defaults p=2, dim=1)


@dataclass(frozen=True)
class MarginConstants:
    """arc_main.py:149-152."""

    s: float
    m: float
    easy_margin: bool
    cos_m: float
    sin_m: float
    th: float
    mm: float


def margin_constants(s: float = 30.0, m: float = 0.50, easy_margin: bool = False) -> MarginConstants:
    return MarginConstants(
        s=float(s),
        m=float(m),
        easy_margin=bool(easy_margin),
        cos_m=math.cos(m),
        sin_m=math.sin(m),
        th=math.cos(math.pi - m),
        mm=math.sin(math.pi - m) * m,
    )


def l2_normalize_rows(a: np.ndarray, eps: float = NORMALIZE_EPS):
    """F.normalize(a) with p=2, dim=1 (arc_main.py:159).  Returns (a_hat, norm_clamped)."""
    a = np.asarray(a, dtype=np.float64)
    nrm = np.sqrt((a * a).sum(axis=1, keepdims=True))
    den = np.maximum(nrm, eps)
    return a / den, den


def labels_as_index(label, batch: int) -> np.ndarray:
    """``label.view(-1, 1).long()`` (arc_main.py:169): any numeric dtype, [B] or [B,1]."""
    idx = np.asarray(label).reshape(-1).astype(np.int64)
    if idx.shape[0] != batch:
        raise IndexError(f"label has {idx.shape[0]} entries for a batch of {batch}")
    return idx


def margin_logits(cosine: np.ndarray, idx: np.ndarray, k: MarginConstants):
    """arc_main.py:160-172 on an already computed cosine matrix.

    Returns (logits, dlogit_dcos) where ``dlogit_dcos`` is the element-wise derivative
    of the returned logits w.r.t. ``cosine`` (what autograd propagates at :329).
    """
    B, C = cosine.shape
    if idx.min(initial=0) < 0 or idx.max(initial=0) >= C:
        raise IndexError("label out of range")  # scatter_ raises in the reference (:169)
    rows = np.arange(B)
    ct = cosine[rows, idx]
    one_minus = 1.0 - ct * ct
    clipped = np.clip(one_minus, 0.0, 1.0)
    sine = np.sqrt(clipped)  # :160
    phi = ct * k.cos_m - sine * k.sin_m  # :161
    if k.easy_margin:
        take_phi = ct > 0  # :163
        alt = ct
    else:
        take_phi = ct > k.th  # :165
        alt = ct - k.mm
    tgt = np.where(take_phi, phi, alt)
    logits = cosine.copy()
    logits[rows, idx] = tgt  # :168-171 one-hot blend
    logits *= k.s  # :172

    # d phi / d cos = cos_m + sin_m * cos / sine while (1-cos^2) is strictly inside the clamp.
    # The reference yields NaN when sine == 0 exactly (inf * 0); the CUDA path returns the
    # finite limit cos_m there (documented divergence) and so does the oracle.
    inside = (one_minus > 0.0) & (one_minus < 1.0) & (sine > 0.0)
    with np.errstate(divide="ignore", invalid="ignore"):
        dphi = np.where(inside, k.cos_m + k.sin_m * ct / np.where(sine > 0, sine, 1.0), k.cos_m)
    dtgt = np.where(take_phi, dphi, 1.0)
    dz = np.full_like(cosine, k.s)
    dz[rows, idx] = k.s * dtgt
    return logits, dz


def head_forward(x, w, label, s=30.0, m=0.50, easy_margin=False) -> np.ndarray:
    """``ArcMarginProduct(...).forward(x, label)`` (arc_main.py:157-176) -> [B, C] logits."""
    k = margin_constants(s, m, easy_margin)
    x = np.asarray(x, dtype=np.float64)
    w = np.asarray(w, dtype=np.float64)
    if x.ndim != 2:
        raise IndexError("input must be 2-D [B, in_features]")  # normalize(dim=1) on 1-D raises
    xh, _ = l2_normalize_rows(x)
    wh, _ = l2_normalize_rows(w)
    cosine = xh @ wh.T  # :159 F.linear
    idx = labels_as_index(label, x.shape[0])
    logits, _ = margin_logits(cosine, idx, k)
    return logits


def cross_entropy_mean(logits: np.ndarray, idx: np.ndarray):
    """nn.CrossEntropyLoss() with default mean reduction (arc_main.py:245,:327).

    Returns (loss, lse[B], softmax[B, C]).
    """
    zmax = logits.max(axis=1, keepdims=True)
    ex = np.exp(logits - zmax)
    ssum = ex.sum(axis=1, keepdims=True)
    lse = (zmax + np.log(ssum)).reshape(-1)
    rows = np.arange(logits.shape[0])
    loss = float((lse - logits[rows, idx]).mean())
    return loss, lse, ex / ssum


def head_forward_backward(x, w, label, s=30.0, m=0.50, easy_margin=False, grad_scale=1.0):
    """Forward + closed-form backward of head + CE (arc_main.py:326-329).

    Returns a dict with: logits, loss, lse, target_logit, argmax, top3, dx, dw.
    ``grad_scale`` is the upstream gradient of the scalar loss (1.0 for loss.backward()).
    """
    k = margin_constants(s, m, easy_margin)
    x = np.asarray(x, dtype=np.float64)
    w = np.asarray(w, dtype=np.float64)
    B = x.shape[0]
    xh, xden = l2_normalize_rows(x)
    wh, wden = l2_normalize_rows(w)
    cosine = xh @ wh.T
    idx = labels_as_index(label, B)
    logits, dz = margin_logits(cosine, idx, k)
    loss, lse, p = cross_entropy_mean(logits, idx)
    rows = np.arange(B)

    g = p.copy()
    g[rows, idx] -= 1.0
    g *= grad_scale / B  # mean reduction
    dcos = g * dz
    dxh = dcos @ wh  # [B, D]
    dwh = dcos.T @ xh  # [C, D]

    def normalize_backward(a_raw, a_hat, den, d_hat):
        # y = a / max(||a||, eps).  Above eps: (d - y (y.d)) / ||a||; clamped: d / eps.
        nrm = np.sqrt((a_raw * a_raw).sum(axis=1, keepdims=True))
        proj = (a_hat * d_hat).sum(axis=1, keepdims=True)
        above = nrm > NORMALIZE_EPS
        return np.where(above, (d_hat - a_hat * proj) / den, d_hat / den)

    dx = normalize_backward(x, xh, xden, dxh)
    dw = normalize_backward(w, wh, wden, dwh)

    order = np.argsort(-logits, axis=1, kind="stable")
    return {
        "logits": logits,
        "loss": loss,
        "lse": lse,
        "target_logit": logits[rows, idx],
        "argmax": order[:, 0],
        "top3": order[:, : min(3, logits.shape[1])],
        "dx": dx,
        "dw": dw,
    }


# --------------------------------------------------------------------------------------
# Class-sharded algebra: what each rank contributes and how it is merged.
# Used by the gloo CPU tests as the per-shard engine and as the checker for the merge.
# --------------------------------------------------------------------------------------

def shard_forward_stats(x, w_shard, label, class_offset, s=30.0, m=0.50, easy_margin=False):
    """Per-shard record for rows of ``x``: (row max, sum-exp relative to it, target logit or -inf)."""
    k = margin_constants(s, m, easy_margin)
    x = np.asarray(x, dtype=np.float64)
    w_shard = np.asarray(w_shard, dtype=np.float64)
    B = x.shape[0]
    Cs = w_shard.shape[0]
    idx = labels_as_index(label, B)
    xh, _ = l2_normalize_rows(x)
    wh, _ = l2_normalize_rows(w_shard)
    cosine = xh @ wh.T
    local = idx - class_offset
    owned = (local >= 0) & (local < Cs)
    logits = cosine * k.s
    tgt = np.full(B, -np.inf)
    if owned.any():
        sub, _ = margin_logits(cosine[owned], local[owned], k)
        logits[owned] = sub
        tgt[owned] = sub[np.arange(int(owned.sum())), local[owned]]
    mx = logits.max(axis=1)
    se = np.exp(logits - mx[:, None]).sum(axis=1)
    return mx, se, tgt


def _normalize_backward(a_raw, d_hat):
    """Backward of F.normalize (arc_main.py:159) for rows ``a_raw`` given d loss / d a_hat."""
    a_raw = np.asarray(a_raw, dtype=np.float64)
    a_hat, den = l2_normalize_rows(a_raw)
    nrm = np.sqrt((a_raw * a_raw).sum(axis=1, keepdims=True))
    proj = (a_hat * d_hat).sum(axis=1, keepdims=True)
    return np.where(nrm > NORMALIZE_EPS, (d_hat - a_hat * proj) / den, d_hat / den)


def shard_backward(x, w_shard, label, class_offset, lse, s=30.0, m=0.50, easy_margin=False, grad_scale=1.0):
    """One class shard's share of the backward, given the GLOBAL log-sum-exp per row.

    Returns (dw_shard [Cs, D] -- final, d loss / d x_hat partial [B, D] -- to be summed over shards).
    """
    k = margin_constants(s, m, easy_margin)
    x = np.asarray(x, dtype=np.float64)
    w_shard = np.asarray(w_shard, dtype=np.float64)
    B, Cs = x.shape[0], w_shard.shape[0]
    idx = labels_as_index(label, B)
    xh, _ = l2_normalize_rows(x)
    wh, _ = l2_normalize_rows(w_shard)
    cosine = xh @ wh.T
    local = idx - class_offset
    owned = (local >= 0) & (local < Cs)
    logits = cosine * k.s
    dz = np.full_like(cosine, k.s)
    if owned.any():
        sub, dsub = margin_logits(cosine[owned], local[owned], k)
        logits[owned] = sub
        dz[owned] = dsub
    g = np.exp(logits - np.asarray(lse, dtype=np.float64)[:, None])
    g[np.nonzero(owned)[0], local[owned]] -= 1.0
    dcos = g * (grad_scale / B) * dz
    return _normalize_backward(w_shard, dcos.T @ xh), dcos @ wh


def input_grad(x, dxh):
    """d loss / d x from the summed d loss / d x_hat (normalisation backward of the embeddings)."""
    return _normalize_backward(x, np.asarray(dxh, dtype=np.float64))


def merge_shard_stats(mx, se, tgt):
    """Merge [P, B] shard records -> (lse[B], target_logit[B], loss)."""
    mx = np.asarray(mx, dtype=np.float64)
    se = np.asarray(se, dtype=np.float64)
    M = mx.max(axis=0)
    S = (se * np.exp(mx - M[None, :])).sum(axis=0)
    lse = M + np.log(S)
    t = np.asarray(tgt, dtype=np.float64).max(axis=0)
    return lse, t, float((lse - t).mean())


# --------------------------------------------------------------------------------------
# Timed CPU baseline: float32 torch port that makes the same dense B x C passes as the
# reference (normalise, SGEMM, margin maths over the whole matrix, one-hot blend, scale,
# CE) and lets autograd derive the backward, like arc_main.py:326-329 does.
# --------------------------------------------------------------------------------------

class TorchPort:
    """fp32 torch-CPU port of head + loss, multi-threaded through ATen (cpu_baseline 'port')."""

    def __init__(self, weight, s=30.0, m=0.50, easy_margin=False):
        import torch

        self.torch = torch
        self.k = margin_constants(s, m, easy_margin)
        self.weight = weight.detach().clone().float().requires_grad_(True)

    def step(self, x, label):
        """One fwd+bwd; returns (loss value, dx, dw)."""
        torch = self.torch
        F = torch.nn.functional
        k = self.k
        x = x.detach().clone().float().requires_grad_(True)
        self.weight.grad = None
        cos = F.linear(F.normalize(x), F.normalize(self.weight))
        sin = (1.0 - cos.pow(2)).clamp(0, 1).sqrt()
        phi = cos * k.cos_m - sin * k.sin_m
        if k.easy_margin:
            phi = torch.where(cos > 0, phi, cos)
        else:
            phi = torch.where(cos > k.th, phi, cos - k.mm)
        hot = torch.zeros_like(cos)
        hot.scatter_(1, label.view(-1, 1).long(), 1)
        out = hot * phi + (1.0 - hot) * cos
        out = out * k.s
        loss = F.cross_entropy(out, label.view(-1).long())
        loss.backward()
        self.last_logits = out.detach()
        return float(loss.detach()), x.grad, self.weight.grad
