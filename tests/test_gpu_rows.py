"""Row kernels through the C ABI against torch: F.normalize (arc_main.py:159) and its backward,
and the accuracy counters of the trainer (arc_main.py:338-345) on the fused outputs."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib
from oracle import arcface_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("D", [8, 100, 512, 2048, 2500])  # register paths of every width + the generic path
def test_l2norm_rows_vs_f_normalize(D):
    rows = 333
    g = torch.Generator().manual_seed(D)
    x = torch.randn(rows, D, generator=g) * torch.logspace(-3, 3, rows).unsqueeze(1)  # norms over six decades
    x[5] = 0.0            # zero row: F.normalize clamps the norm at eps = 1e-12 -> a zero unit row, 1/norm = 1e12
    x[6] = 1e-20          # norm below eps
    xd = x.cuda()
    L = _lib.lib()
    Dp = L.arcb_padded_dim(D)
    out = torch.full((rows, Dp), float("nan"), dtype=torch.bfloat16, device="cuda")
    inv = torch.full((rows,), float("nan"), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(L.arcb_l2norm_rows(xd.data_ptr(), rows, D, D, out.data_ptr(), Dp, inv.data_ptr(), st), "arcb_l2norm_rows")
    _lib.check(L.arcb_check_watchdog(st))
    ref = F.normalize(xd.double(), dim=1)  # p = 2, eps = 1e-12: the reference call
    ref_inv = 1.0 / torch.linalg.vector_norm(xd.double(), dim=1).clamp_min(1e-12)
    # unit rows: bf16 rounding of an fp32 result (relative 2^-9 per element) -- the tolerance SURVEY.md 7.2 states
    got = out[:, :D].double()
    assert torch.all((got - ref).abs() <= ref.abs() * 2.0 ** -8 + 1e-30)
    assert torch.all(out[:, D:] == 0)                                   # zero padding up to Dp (GEMM operand)
    # inv_norm: fp32, relative 1e-6
    torch.testing.assert_close(inv.double(), ref_inv, rtol=1e-6, atol=0.0)
    assert inv[5].item() == pytest.approx(1e12, rel=1e-6) and torch.all(out[5] == 0)


@pytest.mark.parametrize("D", [8, 100, 512, 2500])
def test_l2norm_rows_bwd_vs_autograd(D):
    rows = 77
    g = torch.Generator().manual_seed(100 + D)
    x = torch.randn(rows, D, generator=g).cuda()
    gy = torch.randn(rows, D, generator=g).cuda()
    L = _lib.lib()
    Dp = L.arcb_padded_dim(D)
    gpad = torch.zeros(rows, Dp, device="cuda")
    gpad[:, :D] = gy
    dx = torch.empty(rows, D, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(L.arcb_l2norm_rows_bwd(x.data_ptr(), D, gpad.data_ptr(), Dp, rows, D, dx.data_ptr(), D, st))
    xr = x.double().requires_grad_(True)
    F.normalize(xr, dim=1).backward(gy.double())
    torch.testing.assert_close(dx.double(), xr.grad, rtol=2e-5, atol=1e-6)


def test_correct_counts_on_gpu_equal_reference_counters():
    """torch.max(out, 1) / out.topk(3) + eq(labels) of arc_main.py:338-345, from the fused outputs (no logits)."""
    B, D, C = 384, 128, 5000
    g = torch.Generator().manual_seed(21)
    w = (torch.rand(C, D, generator=g) * 2 - 1) * 0.05
    x = torch.randn(B, D, generator=g)
    y = torch.randint(0, C, (B,), generator=g)
    x[: B // 2] = w[y[: B // 2]] * 40 + 0.3 * x[: B // 2]   # half of the rows sit near their class: hits to count
    head = ArcMarginProduct(D, C, s=30.0, m=0.5, easy_margin=True).cuda()
    with torch.no_grad():
        head.weight.copy_(w)
        head.forward_loss(x.cuda(), y.cuda(), want_top3=True)
    counts = head.correct_counts(y.cuda())
    logits = torch.tensor(orc.head_forward(x.numpy(), w.numpy(), y.numpy(), 30.0, 0.5, True))
    top = logits.topk(3, 1, True, True).indices                       # arc_main.py:342
    want = [int((top[:, 0] == y).sum()), int((top == y[:, None]).any(1).sum()), B]
    got = counts.tolist()
    assert got[2] == B and want[0] > B // 4
    # identical wherever the fp32 reference is unambiguous; a tie below fp32 resolution may move one count
    assert abs(got[0] - want[0]) <= 1 and abs(got[1] - want[1]) <= 1
    assert head.correct_counts().tolist() == got                       # labels kept from the forward
