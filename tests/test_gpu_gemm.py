"""tcgen05 GEMM core (csrc/gemm_sm100.cuh) through the C ABI vs torch fp32 matmul on the same
bf16-rounded operands.  Covers every operand-major combination, ragged M/N/K and split-K."""
import pytest
import torch

from ddp_classification_pytorch_b200 import _lib

pytestmark = pytest.mark.gpu


def _run(M, N, K, a_mn, b_mn, split_k, seed=0, pair=False, wide=False):
    dev = torch.device("cuda:0")
    g = torch.Generator(device="cpu").manual_seed(seed)
    A = torch.randn(M, K, generator=g).to(dev).bfloat16()
    B = torch.randn(N, K, generator=g).to(dev).bfloat16()
    ref = A.float() @ B.float().t()
    def pad8(t):  # leading dimension: multiple of 64 elements (covers the 16-byte rule and whole MN blocks)
        r, c = t.shape
        cp = (c + 63) // 64 * 64
        buf = torch.zeros(r, cp, dtype=t.dtype, device=dev)
        buf[:, :c] = t
        return buf
    a_store = pad8(A.t().contiguous()) if a_mn else pad8(A)
    b_store = pad8(B.t().contiguous()) if b_mn else pad8(B)
    out = torch.zeros(M, N, dtype=torch.float32, device=dev)
    L = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(L.arcb_gemm_bf16(a_store.data_ptr(), int(a_mn), b_store.data_ptr(), int(b_mn), M, N, K,
                                a_store.shape[1], b_store.shape[1], out.data_ptr(), N, split_k,
                                (_lib.GEMM_PAIR if pair else 0) | (_lib.GEMM_PAIR_WIDE if wide else 0), st))
    _lib.check(L.arcb_check_watchdog(st))
    err = (out - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    return err / scale


@pytest.mark.parametrize("pair", [False, True])  # one CTA per tile (cta_group::1) / CTA pairs (cta_group::2)
@pytest.mark.parametrize("a_mn", [False, True])
@pytest.mark.parametrize("b_mn", [False, True])
@pytest.mark.parametrize("shape", [(128, 256, 64), (128, 256, 512), (256, 512, 192), (100, 300, 72), (1024, 777, 512)])
def test_gemm_majors(a_mn, b_mn, shape, pair):
    M, N, K = shape
    assert _run(M, N, K, a_mn, b_mn, 1, pair=pair) < 2e-3


@pytest.mark.parametrize("pair", [False, True])
@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (True, True)])
def test_gemm_split_k(a_mn, b_mn, pair):
    assert _run(256, 512, 64 * 37, a_mn, b_mn, 5, pair=pair) < 2e-3


def test_gemm_many_tiles_persistent():
    # more tiles than SMs: exercises the smem ring wrap-around and both TMEM accumulator stages
    assert _run(1024, 256 * 40, 128, False, False, 1) < 2e-3


@pytest.mark.parametrize("shape,split_k", [((256, 512, 64), 1), ((256, 512, 64 * 37), 5), ((1024, 512, 64 * 50), 18),
                                           ((520, 320, 1000), 3), ((100, 777, 200), 1), ((512, 1024, 64 * 9), 1)])
def test_gemm_pair_wide_tiles(shape, split_k):
    """CTA pairs computing TWO 256-column tiles from one A stage (both accumulators = all of TMEM, no double buffering): the
    dX GEMM of the stored-probabilities backward.  Ragged second column tile, split-K, several tiles per pair in a row."""
    M, N, K = shape
    assert _run(M, N, K, True, True, split_k, pair=True, wide=True) < 2e-3
