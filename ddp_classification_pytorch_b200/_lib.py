"""ctypes binding of ``libarcface_b200.so`` (the C ABI in ``include/arcface_b200.h``).

The shared library is built in-tree by :func:`build` (``nvcc`` for sm_100a) and lives next to
this file so that it travels with the repository snapshot.  There is deliberately no fallback:
if the library is missing or a call fails, a ``RuntimeError`` is raised.
"""
from __future__ import annotations

import ctypes
import os
import shutil
import subprocess
import threading

_PKG_DIR = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_PKG_DIR)
LIB_PATH = os.path.join(_PKG_DIR, "libarcface_b200.so")
_SRC = os.path.join(_PKG_DIR, "csrc", "arcface_b200.cu")
_HDRS = [
    os.path.join(_PKG_DIR, "csrc", n)
    for n in ("ptx_sm100.cuh", "gemm_sm100.cuh", "gemm2_sm100.cuh", "epilogues.cuh", "rowwise.cuh", "bwd_persistent.cuh",
              "optim_rows.cuh", "fwd_resident.cuh", "bwd_probs.cuh")
] + [os.path.join(_ROOT, "include", "arcface_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC",
]

# every symbol include/arcface_b200.h declares (checked by the CPU test-suite)
EXPORTED = [
    "arcb_version", "arcb_last_error", "arcb_check_watchdog", "arcb_l2norm_rows", "arcb_l2norm_rows_bwd",
    "arcb_gemm_bf16", "arcb_ws_bytes", "arcb_scratch_bytes", "arcb_shard_forward", "arcb_shard_logits",
    "arcb_merge_loss", "arcb_shard_backward", "arcb_padded_dim", "arcb_launch_count", "arcb_profile_enable",
    "arcb_profile_collect", "arcb_shard_forward_static", "arcb_shard_top3",
    "arcb_shard_forward_prenorm", "arcb_ws_weight_slots", "arcb_sgd_step", "arcb_adam_step",
    "arcb_suggest_chunk_classes", "arcb_shard_forward_peer", "arcb_merge_loss_peer", "arcb_dx_peer_reduce",
    "arcb_keep_probs_active",
]

MAX_PEERS = 16


class PeerExchange(ctypes.Structure):
    """arcb_peer_exchange of include/arcface_b200.h."""
    _fields_ = [("rank", ctypes.c_int), ("world", ctypes.c_int), ("epoch", ctypes.c_uint),
                ("gather_bufs", ctypes.c_void_p * MAX_PEERS), ("flags", ctypes.c_void_p * MAX_PEERS)]


class PeerReduce(ctypes.Structure):
    """arcb_peer_reduce of include/arcface_b200.h."""
    _fields_ = [("rank", ctypes.c_int), ("world", ctypes.c_int), ("epoch", ctypes.c_uint), ("own_only", ctypes.c_int),
                ("partials", ctypes.c_void_p * MAX_PEERS), ("slices", ctypes.c_void_p * MAX_PEERS),
                ("ready", ctypes.c_void_p * MAX_PEERS), ("done", ctypes.c_void_p * MAX_PEERS),
                ("counter", ctypes.c_void_p)]

# per-call flags of include/arcface_b200.h (0 = library defaults)
FWD_NORM_SEPARATE = 0x1
FWD_NORM_FUSED = 0x2
FWD_PAIR = 0x4
FWD_STREAMING = 0x8
FWD_BLOCK_MAXIMA = 0x200
GEMM_PAIR = 0x4
GEMM_PAIR_WIDE = 0x8
BWD_PER_CHUNK = 0x10
BWD_SINGLE_CTA = 0x20
BWD_STATIC_SHARE = 0x40
BWD_NO_COOP = 0x80  # profiling only
BWD_RECOMPUTE = 0x100

KERNEL_KINDS = [
    "l2norm_x", "l2norm_w", "target_prep", "fwd_gemm", "merge", "loss", "bwd_prep", "bwd_g_gemm", "dx_gemm",
    "dw_gemm", "l2norm_bwd", "logits_gemm", "gemm_test", "bwd_persistent", "top3", "optim",
]

_lock = threading.Lock()
_lib = None


def _nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: cannot build libarcface_b200.so")
    return exe


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    built = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(p) > built for p in [_SRC] + _HDRS)


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the CUDA extension for sm_100a in-tree.  Returns the library path."""
    if not force and not needs_build():
        return LIB_PATH
    cmd = [_nvcc()] + NVCC_FLAGS + ["-I", os.path.join(_ROOT, "include"), "-o", LIB_PATH, _SRC]
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB_PATH


def _declare(lib):
    c = ctypes
    vp, i64, i32, f32, f64 = c.c_void_p, c.c_int64, c.c_int, c.c_float, c.c_double
    lib.arcb_version.restype = i32
    lib.arcb_version.argtypes = []
    lib.arcb_last_error.restype = c.c_char_p
    lib.arcb_last_error.argtypes = []
    lib.arcb_check_watchdog.restype = i32
    lib.arcb_check_watchdog.argtypes = [vp]
    lib.arcb_padded_dim.restype = i64
    lib.arcb_padded_dim.argtypes = [i64]
    lib.arcb_l2norm_rows.restype = i32
    lib.arcb_l2norm_rows.argtypes = [vp, i64, i64, i64, vp, i64, vp, vp]
    lib.arcb_l2norm_rows_bwd.restype = i32
    lib.arcb_l2norm_rows_bwd.argtypes = [vp, i64, vp, i64, i64, i64, vp, i64, vp]
    lib.arcb_gemm_bf16.restype = i32
    lib.arcb_gemm_bf16.argtypes = [vp, i32, vp, i32, i64, i64, i64, i64, i64, vp, i64, i32, i32, vp]
    lib.arcb_ws_bytes.restype = i64
    lib.arcb_ws_bytes.argtypes = [i64, i64, i64, i32]
    lib.arcb_suggest_chunk_classes.restype = i64
    lib.arcb_suggest_chunk_classes.argtypes = [i64, i64, i64]
    lib.arcb_keep_probs_active.restype = i32
    lib.arcb_keep_probs_active.argtypes = [i64, i64, i64, f32, i32]
    lib.arcb_scratch_bytes.restype = i64
    lib.arcb_scratch_bytes.argtypes = [i64, i64]
    fwd_args = [vp, i64, vp, i64, vp, i64, i64, i64, i64, f32, f32, i32, vp, i64]
    lib.arcb_shard_forward.restype = i32
    lib.arcb_shard_forward.argtypes = fwd_args + [vp, i32, vp]
    lib.arcb_shard_forward_prenorm.restype = i32
    lib.arcb_shard_forward_prenorm.argtypes = fwd_args + [vp, i32, vp]
    lib.arcb_shard_forward_peer.restype = i32
    lib.arcb_shard_forward_peer.argtypes = fwd_args + [vp, i32, c.POINTER(PeerExchange), vp]
    lib.arcb_merge_loss_peer.restype = i32
    lib.arcb_merge_loss_peer.argtypes = [vp, vp, c.c_uint, i64, i64, vp, vp, vp, vp]
    lib.arcb_dx_peer_reduce.restype = i32
    lib.arcb_dx_peer_reduce.argtypes = [vp, i64, i64, i64, c.POINTER(PeerReduce), vp, i64, vp]
    lib.arcb_ws_weight_slots.restype = i32
    lib.arcb_ws_weight_slots.argtypes = [vp, i64, i64, c.POINTER(vp), c.POINTER(vp)]
    lib.arcb_sgd_step.restype = i32
    lib.arcb_sgd_step.argtypes = [vp, i64, vp, i64, vp, i64, i64, i64, f64, f64, f64, f64, i32, i32, vp, i64, vp, vp]
    lib.arcb_adam_step.restype = i32
    lib.arcb_adam_step.argtypes = [vp, i64, vp, i64, vp, vp, i64, i64, i64, f64, f64, f64, f64, f64, i64, vp, i64, vp, vp]
    lib.arcb_shard_logits.restype = i32
    lib.arcb_shard_logits.argtypes = fwd_args + [vp, i64, vp]
    lib.arcb_shard_forward_static.restype = i32
    lib.arcb_shard_forward_static.argtypes = fwd_args[:-2] + [vp, vp]
    lib.arcb_shard_top3.restype = i32
    lib.arcb_shard_top3.argtypes = [vp, i64, vp, i64, i64, i64, i64, i64, f32, vp, i64, vp, vp, i32, vp]
    lib.arcb_merge_loss.restype = i32
    lib.arcb_merge_loss.argtypes = [vp, i64, i64, vp, vp, vp, vp]
    lib.arcb_shard_backward.restype = i32
    lib.arcb_shard_backward.argtypes = [i64, i64, i64, f32, vp, i64, vp, vp, vp, vp, i64, vp, i64, vp, vp, i32, vp]
    lib.arcb_launch_count.restype = i64
    lib.arcb_launch_count.argtypes = []
    lib.arcb_profile_enable.restype = i32
    lib.arcb_profile_enable.argtypes = [i32]
    lib.arcb_profile_collect.restype = i32
    lib.arcb_profile_collect.argtypes = [vp, vp, i32]


def lib():
    """The loaded library.  Raises RuntimeError when it has not been built (no fallback path)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                    "(the ArcFace head has no CPU or eager-PyTorch fallback)")
            handle = ctypes.CDLL(LIB_PATH)
            _declare(handle)
            _lib = handle
    return _lib


def profile_collect():
    """{kind: (milliseconds, launches)} since the last collect (synchronises the device)."""
    n = len(KERNEL_KINDS)
    ms = (ctypes.c_double * n)()
    cnt = (ctypes.c_int64 * n)()
    check(lib().arcb_profile_collect(ctypes.cast(ms, ctypes.c_void_p), ctypes.cast(cnt, ctypes.c_void_p), n),
          "arcb_profile_collect")
    return {KERNEL_KINDS[i]: (ms[i], cnt[i]) for i in range(n)}


def check(rc: int, what: str = "arcface_b200") -> None:
    if rc != 0:
        msg = lib().arcb_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (code {rc}): {msg}")
