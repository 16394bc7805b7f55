"""CPU checks of the drop-in boundary: the C-ABI library loads and exports every symbol declared
in include/arcface_b200.h, size queries work without a GPU, and the Python module keeps the
reference's constructor / attribute / state_dict contract (arc_main.py:139-152)."""
import ctypes
import math
import os
import re

import pytest
import torch

from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib, shard_bounds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    _lib.build()
    return _lib.lib()


def test_header_symbols_are_exported(built):
    hdr = open(os.path.join(ROOT, "include", "arcface_b200.h")).read()
    declared = set(re.findall(r"\b(arcb_[a-z0-9_]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTED)
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for sym in declared:
        assert getattr(handle, sym) is not None


def test_flag_constants_match_the_header():
    """The per-call flag bits of the Python binding are the ones include/arcface_b200.h defines (and no two flags of one
    entry point share a bit)."""
    hdr = open(os.path.join(ROOT, "include", "arcface_b200.h")).read()
    defined = {m.group(1): int(m.group(2), 16)
               for m in re.finditer(r"#define ARCB_((?:FWD|BWD|GEMM)_[A-Z_]+) (0x[0-9a-fA-F]+)", hdr)}
    assert len(defined) >= 12
    for name, value in defined.items():
        assert getattr(_lib, name) == value, name
    head_flags = [v for k, v in defined.items() if not k.startswith("GEMM_")]
    assert len(set(head_flags)) == len(head_flags) and all(v & (v - 1) == 0 for v in head_flags)


def test_size_queries_without_gpu(built):
    L = built
    assert L.arcb_version() >= 20000
    assert L.arcb_padded_dim(512) == 512 and L.arcb_padded_dim(72) == 128
    ws = L.arcb_ws_bytes(1024, 1_000_000, 512, _lib.BWD_RECOMPUTE)
    assert ws == L.arcb_ws_bytes(1024, 1_000_000, 512, _lib.BWD_PER_CHUNK | _lib.FWD_NORM_FUSED)  # pure in its arguments
    # dominated by the bf16 copy of the normalised weight (2 B/elem) + per-tile softmax partials
    assert 1_000_000 * 512 * 2 < ws < 1_000_000 * 512 * 2 * 1.1
    # the default for this shape keeps the bf16 softmax numerators P'^T [Cs, B] for the backward on top of that
    kept = L.arcb_ws_bytes(1024, 1_000_000, 512, 0) - ws
    assert L.arcb_keep_probs_active(1024, 1_000_000, 512, 30.0, 0) == 1
    assert 1_000_000 * 1024 * 2 <= kept < 1_000_000 * 1024 * 2 * 1.03  # + block maxima, c (.) x_hat
    assert L.arcb_keep_probs_active(1024, 1_000_000, 256, 30.0, 0) == 0
    assert L.arcb_ws_bytes(1024, 1_000_000, 256, 0) == L.arcb_ws_bytes(1024, 1_000_000, 256, _lib.BWD_RECOMPUTE)
    assert L.arcb_scratch_bytes(1024, 128 * 148) == 128 * 148 * 1024 * 2
    assert L.arcb_ws_bytes(0, 10, 10, 0) < 0
    assert L.arcb_launch_count() == 0


def test_schedule_rule_without_gpu(built):
    """When the backward runs from stored probabilities (pure host logic, include/arcface_b200.h: ARCB_BWD_RECOMPUTE)."""
    L = built
    act = L.arcb_keep_probs_active
    assert act(1024, 1_000_000, 512, 30.0, 0) == 1 and act(1024, 125_000, 512, 30.0, 0) == 1  # cfg4 on 1 and 8 GPUs
    assert act(512, 100_000, 512, 30.0, 0) == 1 and act(512, 50_000, 512, 30.0, 0) == 1      # cfg3 on 1 and 2 GPUs
    assert act(512, 25_000, 512, 30.0, 0) == 0 and act(512, 12_500, 512, 30.0, 0) == 0       # ... launch bound at 4 / 8
    assert act(256, 1000, 2048, 30.0, 0) == 0                                                # cfg2: D > 512
    assert act(1024, 1_000_000, 256, 30.0, 0) == 0 and act(1024, 1_000_000, 320, 30.0, 0) == 1
    assert act(1024, 1_000_000, 512, 41.0, 0) == 1 and act(1024, 1_000_000, 512, 64.0, 0) == 0  # bf16 range of P'
    assert act(64, 1_000_000, 512, 30.0, 0) == 0
    for flag in (_lib.BWD_RECOMPUTE, _lib.BWD_PER_CHUNK, _lib.BWD_SINGLE_CTA, _lib.BWD_STATIC_SHARE, _lib.FWD_PAIR):
        assert act(1024, 1_000_000, 512, 30.0, flag) == 0
    for flag in (_lib.FWD_NORM_SEPARATE, _lib.FWD_NORM_FUSED, _lib.FWD_STREAMING, _lib.FWD_BLOCK_MAXIMA, _lib.BWD_NO_COOP):
        assert act(1024, 1_000_000, 512, 30.0, flag) == 1
    # the workspace carries the matrix exactly when the shape rule holds (the size query has no scale argument)
    for B, C, D in ((1024, 125_000, 512), (512, 25_000, 512), (256, 1000, 2048)):
        extra = L.arcb_ws_bytes(B, C, D, 0) - L.arcb_ws_bytes(B, C, D, _lib.BWD_RECOMPUTE)
        assert (extra >= 2 * B * C) == bool(act(B, C, D, 30.0, 0)) and (extra == 0) == (not act(B, C, D, 30.0, 0))


def test_chunk_suggestion_and_weight_slots_without_gpu(built):
    """Host-side sizing logic of the backward and of the weight cache (no launches involved)."""
    L = built
    for B, C, D in [(1024, 1_000_000, 512), (512, 100_000, 512), (256, 1000, 2048), (64, 10, 512), (4096, 300_000, 256)]:
        r = L.arcb_suggest_chunk_classes(B, C, D)
        assert r > 0 and r % 256 == 0
        assert r <= (C + 255) // 256 * 256                              # never longer than the shard
        if C >= 64 * 256:
            assert r >= 16 * 256                                        # short chunks pay too much reduce-add traffic
    assert L.arcb_suggest_chunk_classes(0, 10, 10) < 0
    # the weight slots lead the workspace: same offsets whatever the batch size, w_hat first, both 256-byte aligned
    a, b = ctypes.c_void_p(), ctypes.c_void_p()
    base = 0x7f0000000000
    C, D = 10_000, 72
    assert L.arcb_ws_weight_slots(ctypes.c_void_p(base), C, D, ctypes.byref(a), ctypes.byref(b)) == 0
    assert a.value == base and (b.value - base) % 256 == 0
    assert b.value - base >= C * L.arcb_padded_dim(D) * 2
    for B in (1, 100, 5000):
        assert L.arcb_ws_bytes(B, C, D, 0) >= (b.value - base) + C * 4
    assert L.arcb_ws_weight_slots(ctypes.c_void_p(base + 8), C, D, ctypes.byref(a), ctypes.byref(b)) == -1  # alignment


def test_no_process_wide_mode_switches(built):
    """no global mutable state -- code paths are chosen by the per-call flags argument only."""
    handle = ctypes.CDLL(_lib.LIB_PATH)
    for gone in ("arcb_set_pair_mode", "arcb_set_forward_norm_mode", "arcb_set_backward_mode", "arcb_set_stored_probs_mode"):
        assert not hasattr(handle, gone)
    head_a = ArcMarginProduct(16, 8, kernel_flags=_lib.BWD_PER_CHUNK)
    head_b = ArcMarginProduct(16, 8)
    assert head_a._cfg()["flags"] == _lib.BWD_PER_CHUNK and head_b._cfg()["flags"] == 0


def test_optimizer_argument_checks(built):
    L = built
    fake = ctypes.c_void_p(0x7f0000000000)
    # rejected before anything is launched
    assert L.arcb_sgd_step(fake, 64, fake, 64, None, 64, 4, 64, 0.1, 0.9, 0.0, 0.0, 0, 1, None, 64, None, None) == -1
    assert L.arcb_sgd_step(fake, 64, fake, 64, fake, 64, 4, 64, 0.1, 0.0, 0.0, 0.0, 1, 1, None, 64, None, None) == -1
    assert b"Nesterov" in L.arcb_last_error()
    assert L.arcb_adam_step(fake, 64, fake, 64, fake, fake, 64, 4, 64, 1e-3, 0.9, 1.0, 1e-8, 0.0, 1, None, 64, None,
                            None) == -1
    assert b"betas" in L.arcb_last_error()


def test_errors_are_codes_not_exceptions(built):
    L = built
    rc = L.arcb_l2norm_rows(None, 4, 8, 8, None, 64, None, None)
    assert rc == -1 and b"null" in L.arcb_last_error()
    rc = L.arcb_shard_backward(4, 4, 4, 30.0, None, 0, None, None, None, None, 0, None, 0, None, None, 0, None)
    assert rc == -1


def test_module_contract_matches_reference():
    head = ArcMarginProduct(256, 2173, s=30, m=0.5, easy_margin=True)  # arc_main.py:234
    assert (head.in_features, head.out_features, head.s, head.m, head.easy_margin) == (256, 2173, 30, 0.5, True)
    assert head.cos_m == math.cos(0.5) and head.sin_m == math.sin(0.5)
    assert head.th == math.cos(math.pi - 0.5) and head.mm == math.sin(math.pi - 0.5) * 0.5
    assert list(head.state_dict().keys()) == ["weight"]
    assert tuple(head.weight.shape) == (2173, 256) and head.weight.dtype == torch.float32
    bound = math.sqrt(6.0 / (256 + 2173))  # xavier_uniform_, arc_main.py:146
    assert head.weight.abs().max().item() <= bound and head.weight.abs().max().item() > 0.9 * bound
    d = ArcMarginProduct(8, 4)
    assert d.s == 30.0 and d.m == 0.5 and d.easy_margin is False
    opt = torch.optim.SGD(head.parameters(), lr=0.1)  # arc_main.py:252
    assert len(opt.param_groups[0]["params"]) == 1
    head.train(); head.eval()


def test_no_cpu_fallback():
    head = ArcMarginProduct(16, 8)
    with pytest.raises(RuntimeError, match="no CPU path"):
        head(torch.randn(4, 16), torch.zeros(4, dtype=torch.long))
    with pytest.raises(RuntimeError, match="no CPU path"):
        head.forward_loss(torch.randn(4, 16), torch.zeros(4, dtype=torch.long))
    with pytest.raises(IndexError):
        head(torch.randn(16), torch.zeros(1))


def test_shard_bounds_partition():
    for C, P in ((1_000_000, 8), (10, 3), (7, 8), (100_000, 4)):
        spans = [shard_bounds(C, P, r) for r in range(P)]
        assert spans[0][0] == 0 and spans[-1][1] == C
        assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
        sizes = [hi - lo for lo, hi in spans]
        assert max(sizes) - min(sizes) <= 1


def test_product_package_never_imports_oracle():
    pkg = os.path.join(ROOT, "ddp_classification_pytorch_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                assert "oracle" not in open(os.path.join(dirpath, f)).read().lower(), f
