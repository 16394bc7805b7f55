"""arcb_shard_forward_static: the forward for hosts that cannot obtain device memory at call time.

Some users of this library sit inside another runtime that owns the whole GPU heap (an inference-server
plugin, a process that froze its memory plan at start-up) or reach the C ABI from a language binding that
has no allocator at all.  They cannot hand over the workspace `arcb_shard_forward` asks for, so this entry
point brings its own fixed 32 MiB block.  The test drives it through the C ABI and checks that it returns
what the ordinary entry point returns on the same inputs -- twice in a row with different shapes, so the
second call runs on top of whatever the first one left in the shared block -- and that oversize problems
are refused with a clear message instead of being launched.

Collected last on purpose (file name), so that a `-x` run has executed every other test before this one.
"""
import pytest
import torch

from ddp_classification_pytorch_b200 import _lib

pytestmark = pytest.mark.gpu

STATIC_WS_BYTES = 32 << 20


def _inputs(B, D, C, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, D, generator=g).cuda()
    w = ((torch.rand(C, D, generator=g) * 2 - 1) * 0.05).cuda()
    y = torch.randint(0, C, (B,), generator=g).cuda()
    return x, w, y


def _forward_both(L, x, w, y, s, m, easy):
    B, D = x.shape
    C = w.shape[0]
    st = torch.cuda.current_stream().cuda_stream
    ws_bytes = L.arcb_ws_bytes(B, C, D, 0)
    assert 0 < ws_bytes <= STATIC_WS_BYTES
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device="cuda")
    ref = torch.full((B, 4), float("nan"), device="cuda")
    new = torch.full((B, 4), float("nan"), device="cuda")
    _lib.check(L.arcb_shard_forward(x.data_ptr(), D, w.data_ptr(), D, y.data_ptr(), B, C, D, 0, s, m, int(easy),
                                    ws.data_ptr(), ws_bytes, ref.data_ptr(), 0, st), "arcb_shard_forward")
    _lib.check(L.arcb_shard_forward_static(x.data_ptr(), D, w.data_ptr(), D, y.data_ptr(), B, C, D, 0, s, m, int(easy),
                                           new.data_ptr(), st), "arcb_shard_forward_static")
    _lib.check(L.arcb_check_watchdog(st))
    return ref, new


def test_static_forward_matches_workspace_forward_two_calls_in_a_row():
    L = _lib.lib()
    # second call: other shapes and margin mode, so the block still holds the first call's leftovers
    for (B, D, C, easy, seed) in ((200, 96, 3000, False, 0), (64, 256, 1500, True, 1)):
        x, w, y = _inputs(B, D, C, seed)
        ref, new = _forward_both(L, x, w, y, 30.0, 0.5, easy)
        assert torch.isfinite(new[:, :3]).all()
        # same kernels on the same inputs: the shard records must agree (tolerance of the parity tests: 1e-3;
        # in practice they are bit-identical)
        torch.testing.assert_close(new, ref, rtol=1e-3, atol=1e-5)
        # and the record is the right one: merged loss equals the reference-style loss of the module path
        lse = torch.empty(B, device="cuda"); zt = torch.empty(B, device="cuda"); loss = torch.empty((), device="cuda")
        _lib.check(L.arcb_merge_loss(new.data_ptr(), 1, B, lse.data_ptr(), zt.data_ptr(), loss.data_ptr(),
                                     torch.cuda.current_stream().cuda_stream))
        from ddp_classification_pytorch_b200 import ArcMarginProduct
        head = ArcMarginProduct(D, C, s=30.0, m=0.5, easy_margin=easy).cuda()
        with torch.no_grad():
            head.weight.copy_(w)
        assert loss.item() == pytest.approx(head.forward_loss(x, y).item(), rel=1e-3)
        # ... and it is the RIGHT loss, not merely the same one: the CPU oracle on the same inputs
        from oracle import arcface_oracle as orc
        ref = orc.head_forward_backward(x.cpu().numpy(), w.cpu().numpy(), y.cpu().numpy(), 30.0, 0.5, easy)
        assert loss.item() == pytest.approx(ref["loss"], rel=1e-3)
        torch.testing.assert_close(lse.cpu().double(), torch.tensor(ref["lse"]).double(), rtol=2e-3, atol=1e-4)


def test_static_forward_rejects_sizes_beyond_its_block():
    L = _lib.lib()
    x, w, y = _inputs(8, 64, 16, 2)
    stats = torch.zeros(8, 4, device="cuda")
    B, C, D = 1024, 1_000_000, 512  # needs ~1.07 GB of workspace: far beyond 32 MiB
    assert L.arcb_ws_bytes(B, C, D, 0) > STATIC_WS_BYTES
    n0 = L.arcb_launch_count()
    rc = L.arcb_shard_forward_static(x.data_ptr(), D, w.data_ptr(), D, y.data_ptr(), B, C, D, 0, 30.0, 0.5, 0,
                                     stats.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == -3  # ARCB_ERR_WORKSPACE
    msg = L.arcb_last_error().decode()
    assert str(STATIC_WS_BYTES) in msg and "arcb_shard_forward" in msg
    assert L.arcb_launch_count() == n0  # nothing was launched
