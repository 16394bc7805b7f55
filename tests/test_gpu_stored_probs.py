"""Experimental stored-probabilities mode (arcb_set_stored_probs_mode): the forward keeps bf16 probabilities relative
to a per-sample reference and the backward runs without its recompute GEMM.  Same parity bar as the default path."""
import numpy as np
import pytest
import torch

from conftest import cosine_similarity
from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib
from oracle import arcface_oracle as orc

pytestmark = pytest.mark.gpu

CASES = [(300, 192, 3000, 30.0, False), (64, 100, 517, 30.0, True), (520, 512, 9000, 64.0, False),
         (129, 66, 257, 30.0, False), (700, 256, 20000, 30.0, True)]


@pytest.mark.parametrize("case", CASES)
def test_stored_probs_mode_matches_oracle_and_default(case):
    B, D, C, s, easy = case
    gen = torch.Generator().manual_seed(B + C)
    w = (torch.rand(C, D, generator=gen) * 2 - 1) * 0.05
    x = torch.randn(B, D, generator=gen)
    n_plant = B // 8  # confident and anti-correlated rows: both ends of the probability range
    y = torch.randint(0, C, (B,), generator=gen)
    x[:n_plant] = w[y[:n_plant]] * 40 + 0.1 * torch.randn(n_plant, D, generator=gen)
    x[n_plant:2 * n_plant] = -w[y[n_plant:2 * n_plant]] * 40 + 0.1 * torch.randn(n_plant, D, generator=gen)
    ref = orc.head_forward_backward(x.numpy(), w.numpy(), y.numpy(), s, 0.5, easy)
    L = _lib.lib()
    res = {}
    try:
        for mode in (1, 0):
            L.arcb_set_stored_probs_mode(mode)
            head = ArcMarginProduct(D, C, s=s, m=0.5, easy_margin=easy, chunk_classes=512).cuda()
            with torch.no_grad():
                head.weight.copy_(w)
            xg = x.cuda().requires_grad_(True)
            loss = head.forward_loss(xg, y.cuda(), want_top3=True)
            loss.backward()
            _lib.check(L.arcb_check_watchdog(torch.cuda.current_stream().cuda_stream))
            res[mode] = (loss.item(), xg.grad.cpu().numpy(), head.weight.grad.cpu().numpy(),
                         head.aux["argmax"].cpu().numpy(), head.aux["lse"].cpu().numpy())
    finally:
        L.arcb_set_stored_probs_mode(0)
    loss, dx, dw, amax, lse = res[1]
    assert abs(loss - ref["loss"]) <= 1e-3 * abs(ref["loss"])
    assert cosine_similarity(dx, ref["dx"]) >= 0.999
    assert cosine_similarity(dw, ref["dw"]) >= 0.999
    assert np.linalg.norm(dw) == pytest.approx(np.linalg.norm(ref["dw"]), rel=3e-2)
    assert not np.isnan(dw).any() and not np.isnan(dx).any()
    # log-sum-exp: bf16 operands dominate the distance to fp64 (s * 2^-9-ish); the stored, rounded probabilities
    # add at most 2^-9 relative to the sum on top of the default path
    np.testing.assert_allclose(lse, ref["lse"], rtol=0, atol=5e-2 * s / 64)
    np.testing.assert_allclose(lse, res[0][4], rtol=0, atol=6e-3)
    # the exact top-1 search works from the same partial format
    assert (amax == res[0][3]).all()
    # and the two modes agree with each other far inside the bar
    assert cosine_similarity(dw, res[0][2]) >= 0.9999 and cosine_similarity(dx, res[0][1]) >= 0.9999
