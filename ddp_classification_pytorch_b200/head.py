"""Drop-in ``ArcMarginProduct`` backed by the sm_100a kernels in ``libarcface_b200.so``.

Mirrors the reference class at ARCFACE/arc_main.py:130-176: same constructor
``(in_features, out_features, s=30.0, m=0.50, easy_margin=False)``, same attributes, same
``forward(input, label)`` contract (returns the ``[B, out_features]`` fp32 scaled logits,
differentiable w.r.t. ``input`` and ``weight``).  Added, non-breaking surface:

* ``forward_loss(input, label)`` -- head + ``nn.CrossEntropyLoss()`` (arc_main.py:245,:327) fused;
  the B x C logits are never written.  Statistics of the last call are kept in ``.aux``.
* ``forward(input, label, fused_loss=True)`` -- the same fused loss reached through ``__call__``, i.e. through a
  ``DistributedDataParallel`` wrapper (arc_main.py:243), so that DDP's reducer all-reduces ``weight.grad``.
* keyword-only ``process_group`` -- class-sharded mode: this rank owns a contiguous slice of the
  classes and only per-row softmax statistics (forward) and the ``[B, D]`` partial of dX
  (backward) cross NVLink through ``torch.distributed``.  ``forward_loss(..., gather_batch=True)`` adds the
  feature / label ``all_gather`` in front and returns dX by ``reduce_scatter`` (data-parallel backbone).

PyTorch is used here for device memory, streams and collectives only; all arithmetic runs in the
hand-written CUDA kernels, reached through the C ABI with raw pointers.  There is no CPU path:
the only engine shipped is :class:`CudaEngine`, which refuses CPU tensors and a missing library.
(The autograd functions talk to the kernels through that small engine interface so that the
collective plumbing can be exercised by CPU ``gloo`` tests with a checker engine of their own.)
"""
from __future__ import annotations

import functools
import math
import os
from typing import Optional

import torch
import torch.nn as nn
from torch.nn import Parameter

from . import _lib

# classes per backward chunk: the persistent backward keeps FOUR chunks of bf16 dLoss/dcos in flight, which
# together with the matching w_hat rows must stay inside the 126 MB L2.  0 = ask the library
# (arcb_suggest_chunk_classes: e.g. 256 * 19 at B=1024, D=512); ARCB_CHUNK_CLASSES overrides it for tuning runs.
DEFAULT_CHUNK_CLASSES = int(os.environ.get("ARCB_CHUNK_CLASSES", 0))
_SCRATCH_BUFS = int(os.environ.get("ARCB_SCRATCH_BUFS", 4))


@functools.lru_cache(maxsize=256)
def _ws_bytes(B, Cs, D, flags):
    return _lib.lib().arcb_ws_bytes(B, Cs, D, flags)


@functools.lru_cache(maxsize=256)
def _scratch_bytes_for(B, Cs, D, chunk_classes, nbuf):
    """Pure size arithmetic of the backward scratch (cached: it sits on the per-step host path)."""
    L = _lib.lib()
    if chunk_classes <= 0:
        chunk_classes = L.arcb_suggest_chunk_classes(B, Cs, D)
    chunk = min(max(128, chunk_classes), (Cs + 127) // 128 * 128)
    return L.arcb_scratch_bytes(B, nbuf * chunk)


@functools.lru_cache(maxsize=64)
def _padded_dim(D):
    return _lib.lib().arcb_padded_dim(D)


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _canonical_inputs(x: torch.Tensor, w: torch.Tensor, label: torch.Tensor):
    """Shape / dtype normalisation shared by every engine (label handling as arc_main.py:169)."""
    if x.dim() != 2:
        raise IndexError(f"input must be 2-D [B, in_features], got {tuple(x.shape)}")  # F.normalize(dim=1)
    if x.shape[1] != w.shape[1]:
        raise RuntimeError(f"input has {x.shape[1]} features, weight expects {w.shape[1]}")
    xc = x.detach()
    if xc.dtype != torch.float32:
        xc = xc.float()
    xc = xc.contiguous()
    wc = w.detach()
    if wc.dtype != torch.float32:
        wc = wc.float()
    wc = wc.contiguous()
    lab = label.detach().reshape(-1)  # label.view(-1, 1).long()
    if lab.dtype != torch.int64:
        lab = lab.long()
    lab = lab.to(x.device).contiguous()
    if lab.numel() != x.shape[0]:
        raise IndexError(f"label has {lab.numel()} entries for a batch of {x.shape[0]}")
    return xc, wc, lab


class _WeightCache:
    """A forward workspace that outlives the call, so that the normalised weight in it can be reused.

    ``F.normalize(self.weight)`` (arc_main.py:159) is recomputed by the reference on every call.  With this
    cache the head keeps ONE workspace; its weight slots (w_hat bf16 + 1/norm) stay valid while the parameter
    is unchanged -- ``valid`` holds the (data_ptr, tensor version) they were computed from -- and the fused
    optimizer step (optim.py) refreshes them in the same pass that updates the weight.  A workspace that an
    un-backpropagated graph still needs is ``busy``: the next forward then takes a fresh, uncached workspace.
    """

    def __init__(self):
        self.ws = None
        self.ws_bytes = 0
        self.key = None
        self.valid = None
        self.busy = 0
        self.generation = 0

    def workspace(self, B, Cs, D, device, flags=0):
        need = _ws_bytes(B, Cs, D, int(flags))
        if need <= 0:
            raise RuntimeError("arcb_ws_bytes rejected the sizes")
        if self.ws is None or self.key != (Cs, D, device) or need > self.ws_bytes:
            self.ws = torch.empty(need, dtype=torch.uint8, device=device)
            self.ws_bytes, self.key, self.valid = need, (Cs, D, device), None
        return self.ws, self.ws_bytes

    def slots(self):
        """Device addresses (w_hat bf16, inv_norm fp32) inside the cached workspace."""
        import ctypes
        Cs, D, _ = self.key
        a, b = ctypes.c_void_p(), ctypes.c_void_p()
        _lib.check(_lib.lib().arcb_ws_weight_slots(self.ws.data_ptr(), Cs, D, ctypes.byref(a), ctypes.byref(b)),
                   "arcb_ws_weight_slots")
        return a.value, b.value


class _Lease:
    """Marks a cached workspace as needed by a pending backward; released by that backward or on collection."""

    def __init__(self, cache):
        self.cache = cache
        self.live = True
        cache.busy += 1

    def release(self):
        if self.live:
            self.live = False
            self.cache.busy -= 1

    __del__ = release


class _PeerGather:
    """Symmetric-memory buffers for the forward's statistics exchange (stage 2).

    Every rank owns ``[2, P, B, 4]`` fp32 (two step slots) plus ``P`` flag words, mapped into every other rank's
    address space (``torch.distributed._symmetric_memory``, NVLink peer access).  The merge kernel of each rank
    stores its ``[B, 4]`` record into all P buffers and raises its flag there; the loss kernel waits for the P flags
    of its own buffer (``arcb_shard_forward_peer`` / ``arcb_merge_loss_peer``).  torch only allocates and maps the
    memory; the exchange itself is done by this package's kernels."""

    def __init__(self, group, B, device):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.P = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.P > _lib.MAX_PEERS:
            raise RuntimeError(f"peer exchange supports up to {_lib.MAX_PEERS} ranks")
        self.B = B
        self.slot_words = self.P * B * 4
        self.buf = symm.empty(2 * self.slot_words + 64, dtype=torch.float32, device=device)
        self.buf.zero_()  # flags start at 0; epochs count from 1
        torch.cuda.synchronize(device)
        self.hdl = symm.rendezvous(self.buf, group)
        self.hdl.barrier()  # every rank's flags are zero before anybody publishes
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.epoch = 0
        self._slots = []  # per step slot: (descriptor, local gathered view); only the epoch changes from step to step
        for slot in range(2):
            px = _lib.PeerExchange()
            px.rank, px.world = self.rank, self.P
            for p in range(self.P):
                px.gather_bufs[p] = self.ptrs[p] + slot * self.slot_words * 4
                px.flags[p] = self.ptrs[p] + 2 * self.slot_words * 4
            gathered = self.buf[slot * self.slot_words:(slot + 1) * self.slot_words].view(self.P * self.B, 4)
            self._slots.append((px, gathered))
        self._flags_ptr = self.ptrs[self.rank] + 2 * self.slot_words * 4

    def next_exchange(self):
        """-> (descriptor for this step, local gathered view [P, B, 4], local flags ptr)."""
        self.epoch += 1
        px, gathered = self._slots[self.epoch & 1]
        px.epoch = self.epoch & 0xFFFFFFFF
        return px, gathered, self._flags_ptr


class _PeerReduce:
    """Symmetric-memory buffers for the backward's dX exchange (stage 2): every rank's ``[B, Dp]`` partial
    of dX_hat (two step slots), its ``[ceil(B / P), D]`` slice of finished rows (two slots), ``P`` ready + ``P`` done flags
    and a block counter.  ``arcb_dx_peer_reduce`` sums the partials over NVLink peer loads, applies F.normalize's backward
    and gathers the rows -- one kernel instead of an NCCL all_reduce plus a row kernel."""

    def __init__(self, group, B, D, Dp, device):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.P = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        if self.P > _lib.MAX_PEERS:
            raise RuntimeError(f"peer exchange supports up to {_lib.MAX_PEERS} ranks")
        self.B, self.D, self.Dp = B, D, Dp
        self.rows_per = (B + self.P - 1) // self.P
        self.part_words = B * Dp
        self.slice_words = self.rows_per * D
        self.flag_off = 2 * self.part_words + 2 * self.slice_words  # then: ready [P], done [P], counter (64-word fields)
        self.buf = symm.empty(self.flag_off + 192, dtype=torch.float32, device=device)
        self.buf.zero_()
        torch.cuda.synchronize(device)
        self.hdl = symm.rendezvous(self.buf, group)
        self.hdl.barrier()
        self.ptrs = [int(p) for p in self.hdl.buffer_ptrs]
        self.epoch = 0
        self._slots = []
        for slot in range(2):
            pr = _lib.PeerReduce()
            pr.rank, pr.world = self.rank, self.P
            for p in range(self.P):
                base = self.ptrs[p]
                pr.partials[p] = base + 4 * slot * self.part_words
                pr.slices[p] = base + 4 * (2 * self.part_words + slot * self.slice_words)
                pr.ready[p] = base + 4 * self.flag_off
                pr.done[p] = base + 4 * (self.flag_off + 64)
            pr.counter = self.ptrs[self.rank] + 4 * (self.flag_off + 128)
            partial = self.buf[slot * self.part_words:(slot + 1) * self.part_words].view(self.B, self.Dp)
            self._slots.append((pr, partial))

    def next_exchange(self, own_only):
        """-> (descriptor of this step, this rank's partial buffer [B, Dp] to hand to the backward kernel)."""
        self.epoch += 1
        pr, partial = self._slots[self.epoch & 1]
        pr.epoch, pr.own_only = self.epoch & 0xFFFFFFFF, int(bool(own_only))
        return pr, partial


class CudaEngine:
    """Every method is one call into libarcface_b200.so (include/arcface_b200.h)."""

    def prepare(self, x, w, label):
        if x.dim() != 2:
            raise IndexError(f"input must be 2-D [B, in_features], got {tuple(x.shape)}")
        if not x.is_cuda or not w.is_cuda:
            raise RuntimeError("ArcMarginProduct (B200) requires CUDA tensors: there is no CPU path")
        _lib.lib()  # raises if the extension has not been built
        return _canonical_inputs(x, w, label)

    @staticmethod
    def _alloc_ws(B, Cs, D, device, flags=0):
        nbytes = _ws_bytes(B, Cs, D, int(flags))
        if nbytes <= 0:
            raise RuntimeError("arcb_ws_bytes rejected the sizes")
        return torch.empty(nbytes, dtype=torch.uint8, device=device), nbytes

    @staticmethod
    def _scratch_for(B, Cs, D, device, chunk_classes):
        nbytes = _scratch_bytes_for(B, Cs, D, int(chunk_classes), _SCRATCH_BUFS)  # four chunk buffers by default
        return torch.empty(nbytes, dtype=torch.uint8, device=device), nbytes

    def shard_forward(self, xc, wc, lab, cfg):
        """-> (stats [B,4], state kept for backward)."""
        L = _lib.lib()
        B, D = xc.shape
        Cs = wc.shape[0]
        stats = torch.empty(B, 4, dtype=torch.float32, device=xc.device)
        cache = cfg.get("wcache")
        tag = cfg.get("w_tag")  # (data_ptr, version) of the parameter itself
        fn, gen = L.arcb_shard_forward, 0
        px = cfg.get("peer_px")  # class-sharded: the merge kernel publishes the record to every rank itself
        if cache is not None and cache.busy == 0 and tag is not None and tag[0] == wc.data_ptr():
            ws, ws_bytes = cache.workspace(B, Cs, D, xc.device, cfg.get("flags", 0))
            if cache.valid == tag and px is None:  # w_hat / inv_norm in the workspace belong to exactly this weight
                fn = L.arcb_shard_forward_prenorm
            cache.valid = None  # until the call below has been enqueued
        else:
            cache = None
            ws, ws_bytes = self._alloc_ws(B, Cs, D, xc.device, cfg.get("flags", 0))
        call_flags = int(cfg.get("flags", 0)) | (_lib.FWD_BLOCK_MAXIMA if cfg.get("want_top3") else 0)
        if px is not None:
            import ctypes
            _lib.check(L.arcb_shard_forward_peer(_ptr(xc), D, _ptr(wc), D, _ptr(lab), B, Cs, D, cfg["class_offset"],
                                                 cfg["s"], cfg["m"], int(cfg["easy_margin"]), _ptr(ws), ws_bytes,
                                                 _ptr(stats), call_flags, ctypes.byref(px), _stream()),
                       "arcb_shard_forward_peer")
        else:
            _lib.check(fn(_ptr(xc), D, _ptr(wc), D, _ptr(lab), B, Cs, D, cfg["class_offset"], cfg["s"], cfg["m"],
                          int(cfg["easy_margin"]), _ptr(ws), ws_bytes, _ptr(stats), call_flags, _stream()),
                       "arcb_shard_forward")
        if cache is not None:
            cache.valid = tag
            cache.generation += 1
            gen = cache.generation
        return stats, (B, Cs, D, ws, ws_bytes, cache, gen)

    def shard_logits(self, xc, wc, lab, cfg):
        L = _lib.lib()
        B, D = xc.shape
        Cs = wc.shape[0]
        ws, ws_bytes = self._alloc_ws(B, Cs, D, xc.device, cfg.get("flags", 0))
        out = torch.empty(B, Cs, dtype=torch.float32, device=xc.device)
        _lib.check(L.arcb_shard_logits(_ptr(xc), D, _ptr(wc), D, _ptr(lab), B, Cs, D, cfg["class_offset"], cfg["s"],
                                       cfg["m"], int(cfg["easy_margin"]), _ptr(ws), ws_bytes, _ptr(out), Cs,
                                       _stream()), "arcb_shard_logits")
        return out, (B, Cs, D, ws, ws_bytes, None, 0)

    def shard_top3(self, xc, wc, state, cfg):
        """-> (values [B,3] fp32 logits, indices [B,3] int64 global class ids) of this shard."""
        L = _lib.lib()
        B, Cs, D, ws, ws_bytes = state[:5]
        val = torch.empty(B, 3, dtype=torch.float32, device=xc.device)
        idx = torch.empty(B, 3, dtype=torch.int64, device=xc.device)
        _lib.check(L.arcb_shard_top3(_ptr(xc), D, _ptr(wc), D, B, Cs, D, cfg["class_offset"], cfg["s"], _ptr(ws),
                                     ws_bytes, _ptr(val), _ptr(idx), int(cfg.get("flags", 0)), _stream()), "arcb_shard_top3")
        return val, idx

    def merge_loss(self, gathered, P, B):
        """stats [P,B,4] -> (lse [B], target_logit [B], loss scalar)."""
        L = _lib.lib()
        dev = gathered.device
        lse = torch.empty(B, dtype=torch.float32, device=dev)
        zt = torch.empty(B, dtype=torch.float32, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        _lib.check(L.arcb_merge_loss(_ptr(gathered), P, B, _ptr(lse), _ptr(zt), _ptr(loss), _stream()),
                   "arcb_merge_loss")
        return lse, zt, loss

    def merge_loss_peer(self, gathered, flags_ptr, epoch, P, B):
        """Like merge_loss, but the kernel first waits until all P ranks have published this step's record."""
        L = _lib.lib()
        dev = gathered.device
        lse = torch.empty(B, dtype=torch.float32, device=dev)
        zt = torch.empty(B, dtype=torch.float32, device=dev)
        loss = torch.empty((), dtype=torch.float32, device=dev)
        _lib.check(L.arcb_merge_loss_peer(_ptr(gathered), flags_ptr, epoch, P, B, _ptr(lse), _ptr(zt), _ptr(loss), _stream()),
                   "arcb_merge_loss_peer")
        return lse, zt, loss

    def dx_peer_reduce(self, xc, pr, rows):
        """Sum of all ranks' dX_hat partials + F.normalize backward, exchanged by the kernel itself -> dx [rows, D]."""
        import ctypes
        L = _lib.lib()
        B, D = xc.shape
        dx = torch.empty(rows, D, dtype=torch.float32, device=xc.device)
        _lib.check(L.arcb_dx_peer_reduce(_ptr(xc), D, B, D, ctypes.byref(pr), _ptr(dx), D, _stream()), "arcb_dx_peer_reduce")
        return dx

    def shard_backward(self, state, cfg, lse=None, zt=None, grad_loss=None, grad_logits=None, dxhat_out=None):
        """-> (dw [Cs,D], dxhat partial [B,Dp]).  ``dxhat_out``: write the partial there (peer-mapped memory)."""
        L = _lib.lib()
        B, Cs, D, ws, ws_bytes = state[:5]
        cache, gen = state[5], state[6]
        if cache is not None and cache.generation != gen:
            raise RuntimeError("ArcMarginProduct (B200): the cached workspace of this forward was reused by a later "
                               "forward before its backward ran (second backward through a retained graph?); "
                               "construct the head with cache_normalized_weight=False for that pattern")
        dev = ws.device
        Dp = _padded_dim(D)
        if grad_logits is None and L.arcb_keep_probs_active(B, Cs, D, cfg["s"], int(cfg.get("flags", 0))):
            scratch, scratch_bytes = None, 0  # the forward kept the probabilities: nothing is recomputed
        else:
            scratch, scratch_bytes = self._scratch_for(B, Cs, D, dev, cfg["chunk_classes"])
        dw = torch.empty(Cs, D, dtype=torch.float32, device=dev)
        dxhat = dxhat_out if dxhat_out is not None else torch.empty(B, Dp, dtype=torch.float32, device=dev)
        _lib.check(L.arcb_shard_backward(B, Cs, D, cfg["s"], _ptr(ws), ws_bytes, _ptr(lse), _ptr(zt), _ptr(grad_loss),
                                         _ptr(grad_logits), Cs if grad_logits is not None else 0, _ptr(scratch),
                                         scratch_bytes, _ptr(dw), _ptr(dxhat), int(cfg.get("flags", 0)), _stream()),
                   "arcb_shard_backward")
        return dw, dxhat

    def input_grad(self, xc, dxhat):
        L = _lib.lib()
        B, D = xc.shape
        dx = torch.empty(B, D, dtype=torch.float32, device=xc.device)
        _lib.check(L.arcb_l2norm_rows_bwd(_ptr(xc), D, _ptr(dxhat), dxhat.shape[1], B, D, _ptr(dx), D, _stream()),
                   "arcb_l2norm_rows_bwd")
        return dx


def _world(group):
    return 1 if group is None else torch.distributed.get_world_size(group)


class _FusedLossFn(torch.autograd.Function):
    """loss = CrossEntropy(ArcMargin(x, W, label)) without materialising the logits."""

    @staticmethod
    def forward(ctx, x, w, label, cfg):
        eng = cfg["engine"]
        group = cfg["group"]
        P = _world(group)
        gather = bool(cfg.get("gather_batch")) and P > 1
        ctx.local_rows = None
        if gather:
            # data-parallel backbone in front of a class-sharded head (partial-FC style): every rank contributes its
            # own rows; the shard needs the global batch.  Rows of rank r land at [r * b, (r + 1) * b).
            if x.dim() != 2:
                raise IndexError(f"input must be 2-D [B, in_features], got {tuple(x.shape)}")
            b_local = x.shape[0]
            xl = x.detach().to(torch.float32).contiguous()
            ll = label.detach().reshape(-1).long().to(x.device).contiguous()
            if ll.numel() != b_local:
                raise IndexError(f"label has {ll.numel()} entries for a batch of {b_local}")
            xg = torch.empty(P * b_local, xl.shape[1], dtype=xl.dtype, device=xl.device)
            lg = torch.empty(P * b_local, dtype=ll.dtype, device=ll.device)
            torch.distributed.all_gather_into_tensor(xg, xl, group=group)
            torch.distributed.all_gather_into_tensor(lg, ll, group=group)
            ctx.local_rows = (torch.distributed.get_rank(group) * b_local, b_local)
            xc, wc, lab = eng.prepare(xg, w, lg)
        else:
            xc, wc, lab = eng.prepare(x, w, label)
        B = xc.shape[0]
        if cfg.get("wcache") is not None:
            cfg = dict(cfg, w_tag=(w.data_ptr(), w._version))
        if not (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]):
            # no backward will follow (evaluation): the forward need not keep its probabilities for one
            cfg = dict(cfg, flags=int(cfg.get("flags", 0)) | _lib.BWD_RECOMPUTE)
        peer = cfg["peer_factory"](B, xc.device) if (P > 1 and cfg.get("peer_factory") is not None) else None
        if peer is not None:  # the exchange is done by the merge / loss kernels themselves over NVLink peer memory
            px, gathered, flags_ptr = peer.next_exchange()
            cfg = dict(cfg, peer_px=px)
        stats, state = eng.shard_forward(xc, wc, lab, cfg)
        lease = None
        if len(state) > 5 and state[5] is not None and (ctx.needs_input_grad[0] or ctx.needs_input_grad[1]):
            lease = _Lease(state[5])  # the workspace now belongs to this graph until its backward has run
        if peer is not None:
            lse, zt, loss = eng.merge_loss_peer(gathered, flags_ptr, px.epoch, P, B)
        else:
            if P > 1:  # the only forward exchange: one [B, 4] record per rank
                gathered = torch.empty(P * B, stats.shape[1], dtype=stats.dtype, device=stats.device)  # = [P, B, 4]
                torch.distributed.all_gather_into_tensor(gathered, stats, group=group)
            else:
                gathered = stats
            lse, zt, loss = eng.merge_loss(gathered, P, B)
        ctx.save_for_backward(xc)
        ctx.held = (eng, state, lse, zt, cfg, w.dtype, x.dtype)
        ctx.lease = lease
        aux = cfg["aux_out"]
        aux.update(lse=lse, target_logit=zt, stats=stats, labels=lab)
        if cfg["want_top3"]:  # predictions for the accuracy counters (arc_main.py:338-345, :377-388)
            val, idx = eng.shard_top3(xc, wc, state, cfg)
            if P > 1:  # keep the best three of the ranks' candidates (index order on equal logits)
                vals = torch.empty(P * B, 3, dtype=val.dtype, device=val.device)
                idxs = torch.empty(P * B, 3, dtype=idx.dtype, device=idx.device)
                torch.distributed.all_gather_into_tensor(vals, val, group=group)
                torch.distributed.all_gather_into_tensor(idxs, idx, group=group)
                vals = vals.view(P, B, 3).permute(1, 0, 2).reshape(B, 3 * P)
                idxs = idxs.view(P, B, 3).permute(1, 0, 2).reshape(B, 3 * P)
                order = torch.argsort(vals, dim=1, descending=True, stable=True)[:, :3]
                val, idx = torch.gather(vals, 1, order), torch.gather(idxs, 1, order)
            aux.update(top3=idx, top3_logits=val, argmax=idx[:, 0])
        return loss

    @staticmethod
    def backward(ctx, grad_loss):
        (xc,) = ctx.saved_tensors
        eng, state, lse, zt, cfg, w_dtype, x_dtype = ctx.held
        g = grad_loss.detach().to(torch.float32).contiguous()
        group = cfg["group"]
        P = _world(group)
        B, D = xc.shape
        reducer = None
        if P > 1 and cfg.get("peer_reduce_factory") is not None and D % 4 == 0 and D <= 2048:
            reducer = cfg["peer_reduce_factory"](B, D, xc.device)
        if reducer is not None:
            # the partial goes straight into peer-mapped memory; ONE kernel then sums all ranks' partials over NVLink,
            # applies F.normalize's backward and collects the rows (or keeps only this rank's rows): no NCCL call
            own = ctx.local_rows is not None
            pr, partial = reducer.next_exchange(own)
            dw, _ = eng.shard_backward(state, cfg, lse=lse, zt=zt, grad_loss=g, dxhat_out=partial)
            if own and ctx.local_rows != (reducer.rank * reducer.rows_per, reducer.rows_per):
                raise RuntimeError("gather_batch needs the same number of rows on every rank")
            dx = eng.dx_peer_reduce(xc, pr, reducer.rows_per if own else B)
            if own:
                dx = dx * float(P)  # see below: the backbone's DDP reducer averages over the ranks
            if ctx.lease is not None:
                ctx.lease.release()
            return dx.to(x_dtype), dw.to(w_dtype), None, None
        dw, dxhat = eng.shard_backward(state, cfg, lse=lse, zt=zt, grad_loss=g)
        if ctx.local_rows is not None:
            # each rank only needs the gradient of its own rows: reduce_scatter of the class shards' partials
            #  instead of an all_reduce of all B rows, then F.normalize backward on those rows.
            # The rows are scaled by the world size: the backbone's DDP reducer will AVERAGE its gradients over the
            # ranks, and the mean of P per-rank terms that each carry 1/B_global must come out as the global-mean
            # gradient the reference's per-rank mean + DDP average produces (arc_main.py:326-329).
            r0, b_local = ctx.local_rows
            mine = torch.empty(b_local, dxhat.shape[1], dtype=dxhat.dtype, device=dxhat.device)
            torch.distributed.reduce_scatter_tensor(mine, dxhat, group=group)
            dx = eng.input_grad(xc[r0:r0 + b_local], mine)
            dx = dx * float(P)
        else:
            if P > 1:  # the only backward exchange: sum of the class shards' dX_hat partials
                torch.distributed.all_reduce(dxhat, group=group)
            dx = eng.input_grad(xc, dxhat)
        if ctx.lease is not None:
            ctx.lease.release()
        return dx.to(x_dtype), dw.to(w_dtype), None, None


class _LogitsFn(torch.autograd.Function):
    """The literal reference forward: returns the [B, Cs] fp32 logits (arc_main.py:157-176)."""

    @staticmethod
    def forward(ctx, x, w, label, cfg):
        eng = cfg["engine"]
        xc, wc, lab = eng.prepare(x, w, label)
        out, state = eng.shard_logits(xc, wc, lab, cfg)
        ctx.save_for_backward(xc)
        ctx.held = (eng, state, cfg, w.dtype, x.dtype)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (xc,) = ctx.saved_tensors
        eng, state, cfg, w_dtype, x_dtype = ctx.held
        go = grad_out.detach().to(torch.float32).contiguous()
        dw, dxhat = eng.shard_backward(state, cfg, grad_logits=go)
        if _world(cfg["group"]) > 1:
            torch.distributed.all_reduce(dxhat, group=cfg["group"])
        dx = eng.input_grad(xc, dxhat)
        return dx.to(x_dtype), dw.to(w_dtype), None, None


def shard_bounds(num_classes: int, world: int, rank: int):
    """Contiguous class range [lo, hi) owned by ``rank`` (sizes differ by at most one)."""
    base, rem = divmod(num_classes, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class ArcMarginProduct(nn.Module):
    r"""Large-margin arc distance head, cos(theta + m)  --  drop-in for arc_main.py:130-176.

    Args:
        in_features: size of each input sample
        out_features: number of classes (the GLOBAL count, also in class-sharded mode)
        s: norm of input feature (logit scale), must be > 0
        m: additive angular margin
        easy_margin: use the ``cos > 0`` fallback instead of ``th`` / ``mm``
        process_group (keyword-only): shard the classes over this ``torch.distributed`` group;
            ``weight`` then holds only this rank's rows and ``forward`` returns only its columns.
        cache_normalized_weight (keyword-only): see :meth:`cache_normalized_weight`.
        peer_exchange (keyword-only): class-sharded mode only.  ``True`` / ``None`` (default: try): the ``[B, 4]``
            statistics are exchanged by the merge / loss kernels themselves through symmetric (NVLink peer) memory
            instead of an NCCL ``all_gather``; ``None`` falls back to NCCL when symmetric memory cannot be set up,
            ``False`` always uses NCCL.  Same results either way.
        kernel_flags (keyword-only): ``ARCB_*`` bits of include/arcface_b200.h passed with every call of THIS head
            (kernel schedule selection for A/B runs and tests; 0 = library defaults).  Per head, not per process.
    """

    def __init__(self, in_features, out_features, s=30.0, m=0.50, easy_margin=False, *, process_group=None,
                 chunk_classes: Optional[int] = None, cache_normalized_weight: bool = False, kernel_flags: int = 0,
                 peer_exchange: Optional[bool] = None, _engine=None):
        super().__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.s = s
        self.m = m
        self.process_group = process_group
        if process_group is not None:
            world = torch.distributed.get_world_size(process_group)
            rank = torch.distributed.get_rank(process_group)
        else:
            world, rank = 1, 0
        self.class_lo, self.class_hi = shard_bounds(out_features, world, rank)
        self.world_size, self.rank = world, rank
        # xavier_uniform_ bound of the FULL matrix (fan-out = global class count, arc_main.py:145-146), so a
        # sharded head starts from the same distribution as a replicated one.
        rows = self.class_hi - self.class_lo
        self.weight = Parameter(torch.empty(rows, in_features, dtype=torch.float32))
        bound = math.sqrt(6.0 / (in_features + out_features))
        with torch.no_grad():
            self.weight.uniform_(-bound, bound)

        self.easy_margin = easy_margin
        self.cos_m = math.cos(m)
        self.sin_m = math.sin(m)
        self.th = math.cos(math.pi - m)
        self.mm = math.sin(math.pi - m) * m
        self.chunk_classes = chunk_classes or DEFAULT_CHUNK_CLASSES  # 0: chosen per call by the library
        self.kernel_flags = int(kernel_flags)
        self.peer_exchange = peer_exchange
        self._peer = {}        # batch size -> _PeerGather
        self._peer_red = {}    # batch size -> _PeerReduce
        self._peer_off = peer_exchange is False or os.environ.get("ARCB_PEER_EXCHANGE", "1") == "0"
        self.aux = {}
        self._engine = _engine if _engine is not None else CudaEngine()
        self._wcache = _WeightCache() if cache_normalized_weight else None

    def cache_normalized_weight(self, enable: bool = True):
        """Keep the forward workspace between calls so that ``F.normalize(self.weight)`` (arc_main.py:159) is only
        recomputed after the weight has changed (``forward_loss`` path).  Switched on by the fused optimizers in
        ``optim.py``, which refresh the normalised copy while they update the weight; also saves the pass in a
        validation loop (arc_main.py:357-404), where the weight is constant."""
        if enable and self._wcache is None:
            self._wcache = _WeightCache()
        elif not enable:
            self._wcache = None
        return self

    def extra_repr(self):
        return (f"in_features={self.in_features}, out_features={self.out_features}, s={self.s}, m={self.m}, "
                f"easy_margin={self.easy_margin}, classes=[{self.class_lo},{self.class_hi})")

    def _cfg(self):
        if not (float(self.s) > 0.0):
            raise ValueError("ArcMarginProduct (B200): the scale s must be positive")
        return {
            "s": float(self.s), "m": float(self.m), "easy_margin": bool(self.easy_margin),
            "class_offset": int(self.class_lo), "group": self.process_group,
            "chunk_classes": int(self.chunk_classes), "aux_out": self.aux, "engine": self._engine,
            "want_top3": False, "wcache": self._wcache, "flags": int(self.kernel_flags), "gather_batch": False,
            "peer_factory": self._peer_gather if (self.world_size > 1 and not self._peer_off
                                                  and isinstance(self._engine, CudaEngine)) else None,
            "peer_reduce_factory": self._peer_reducer if (self.world_size > 1 and not self._peer_off
                                                          and isinstance(self._engine, CudaEngine)) else None,
        }

    def _peer_reducer(self, B, D, device):
        """The symmetric-memory buffers of the backward's dX exchange for batch size B, or None (-> NCCL)."""
        if self._peer_off:
            return None
        pr = self._peer_red.get(B)
        if pr is None:
            ok = torch.ones(1, device=device)
            try:
                pr = _PeerReduce(self.process_group, B, D, _lib.lib().arcb_padded_dim(D), device)
            except Exception as ex:
                if self.peer_exchange is True:
                    raise
                pr, ok[0] = None, 0.0
                self._peer_error = f"{type(ex).__name__}: {ex}"
            torch.distributed.all_reduce(ok, op=torch.distributed.ReduceOp.MIN, group=self.process_group)
            if ok.item() < 0.5:
                self._peer_off, pr = True, None
            else:
                self._peer_red[B] = pr
        return pr

    def _peer_gather(self, B, device):
        """The symmetric-memory exchange buffers for batch size B (set up collectively on first use), or None."""
        if self._peer_off:
            return None
        pg = self._peer.get(B)
        if pg is None:
            ok = torch.ones(1, device=device)
            try:
                pg = _PeerGather(self.process_group, B, device)
            except Exception as ex:  # no peer access / symmetric memory on this system
                if self.peer_exchange is True:
                    raise
                pg, ok[0] = None, 0.0
                self._peer_error = f"{type(ex).__name__}: {ex}"
            # every rank must take the same path: one rank without symmetric memory switches all of them to NCCL
            torch.distributed.all_reduce(ok, op=torch.distributed.ReduceOp.MIN, group=self.process_group)
            if ok.item() < 0.5:
                self._peer_off, pg = True, None
            else:
                self._peer[B] = pg
        return pg

    def forward(self, input, label, fused_loss=False, want_top3=False, gather_batch=False):
        """Scaled logits ``[B, classes of this rank]`` exactly as arc_main.py:157-176 returns them.

        ``fused_loss=True`` returns ``nn.CrossEntropyLoss()(logits, label)`` instead (see :meth:`forward_loss`) without
        writing the logits.  This is the form to use when the head is wrapped the way the reference wraps it,
        ``ARC = DistributedDataParallel(ARC)`` (arc_main.py:243): ``loss = ARC(features, labels, fused_loss=True)`` goes
        through the wrapper's ``forward``, so DDP's reducer is armed and ``weight.grad`` is all-reduced in backward.
        Calling ``ARC.module.forward_loss(...)`` would bypass the wrapper and leave every replica with its own
        un-reduced gradient.
        """
        if fused_loss:
            return self.forward_loss(input, label, want_top3=want_top3, gather_batch=gather_batch)
        return _LogitsFn.apply(input, self.weight, label, self._cfg())

    def forward_loss(self, input, label, want_top3=False, gather_batch=False):
        """``nn.CrossEntropyLoss()(self(input, label), label)`` fused (arc_main.py:326-327).

        ``want_top3=True`` additionally leaves what the trainer's accuracy counters take from the logits
        (``torch.max(out, 1)``, ``out.topk(3)``, arc_main.py:338-345) in ``.aux``: ``argmax`` [B],
        ``top3`` [B,3] (global class ids) and ``top3_logits`` -- exact fp32 re-scores, not bf16 values.

        ``gather_batch=True`` (class-sharded head behind a data-parallel backbone, arc_main.py:320-330 with the head
        sharded instead of replicated): ``input`` / ``label`` are this rank's rows only (same count on every rank); they
        are all-gathered so that every class shard sees the global batch, the loss is the mean over the GLOBAL batch
        (identical on all ranks; ``aux['labels']`` holds the gathered labels), and backward returns this rank's rows of
        dX via ``reduce_scatter``, multiplied by the world size so that the backbone's DDP gradient AVERAGE equals the
        global-mean gradient.  ``weight.grad`` is the plain gradient of that global-mean loss (what the reference's
        DDP-averaged replicas hold); the head itself must NOT be wrapped in DDP in this mode.

        Not reachable through a DDP wrapper: under ``DistributedDataParallel(ARC)`` call
        ``ARC(input, label, fused_loss=True)`` instead (see :meth:`forward`).
        """
        cfg = self._cfg()
        cfg["want_top3"] = bool(want_top3)
        cfg["gather_batch"] = bool(gather_batch)
        return _FusedLossFn.apply(input, self.weight, label, cfg)

    # ---- accuracy counters of the trainer / validation loop (arc_main.py:335-345, :377-395) -------------------
    @torch.no_grad()
    def correct_counts(self, label=None, reduce_group=None):
        """``(top1_correct, top3_correct, samples)`` as a device int64 tensor ``[3]`` from the ``aux`` of the last
        ``forward_loss(..., want_top3=True)`` -- what ``torch.max(out, 1)`` / ``out.topk(3)`` + ``eq(labels)`` give
        in the reference, without the logits and without a host sync.  ``reduce_group``: sum the counters over
        the ranks of that group (data-parallel validation: every rank scored different samples); the reference
        multiplies rank 0's counts by the world size instead (arc_main.py:393-395).  Do not pass a group in
        class-sharded mode: there every rank already scored the same global batch."""
        if "top3" not in self.aux:
            raise RuntimeError("call forward_loss(input, label, want_top3=True) first")
        top3 = self.aux["top3"]
        if label is None:  # the labels the last forward scored (the gathered ones with gather_batch=True)
            label = self.aux["labels"]
        lab = label.detach().reshape(-1).to(device=top3.device, dtype=top3.dtype)
        hit = top3 == lab[:, None]
        counts = torch.stack([hit[:, 0].sum(), hit.any(dim=1).sum(),
                              torch.tensor(lab.numel(), device=top3.device, dtype=torch.int64)]).to(torch.int64)
        if reduce_group is not None and torch.distributed.get_world_size(reduce_group) > 1:
            torch.distributed.all_reduce(counts, group=reduce_group)
        return counts

    # ---- checkpoints (the reference never saves the head, arc_main.py:407-410 commented out) ------------------
    def full_state_dict(self):
        """``{'weight': [out_features, in_features]}`` -- the reference class's ``state_dict()`` layout, also in
        class-sharded mode (gathers the shards; collective: every rank of the group must call it)."""
        return {"weight": self.full_weight()}

    @torch.no_grad()
    def load_full_state_dict(self, state_dict):
        """Load a full ``[out_features, in_features]`` weight (a reference-layout checkpoint, or one written with
        another world size): every rank keeps the rows of its own class range."""
        w = state_dict["weight"]
        if tuple(w.shape) != (self.out_features, self.in_features):
            raise RuntimeError(f"weight has shape {tuple(w.shape)}, expected {(self.out_features, self.in_features)}")
        self.weight.copy_(w[self.class_lo:self.class_hi].to(device=self.weight.device, dtype=self.weight.dtype))
        return self

    @torch.no_grad()
    def full_weight(self):
        """The complete ``[out_features, in_features]`` weight (gathers the shards) for checkpoints."""
        if self.process_group is None or self.world_size == 1:
            return self.weight.detach().clone()
        sizes = [shard_bounds(self.out_features, self.world_size, r) for r in range(self.world_size)]
        rows = max(hi - lo for lo, hi in sizes)
        pad = torch.zeros(rows, self.in_features, dtype=self.weight.dtype, device=self.weight.device)
        pad[: self.weight.shape[0]] = self.weight
        parts = [torch.empty_like(pad) for _ in range(self.world_size)]
        torch.distributed.all_gather(parts, pad, group=self.process_group)
        return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, sizes)], dim=0)
