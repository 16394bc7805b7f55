"""CUDA path vs the reference's golden vectors and vs the oracle, through the Python module that
wraps the C ABI.  Tolerances are the ones BASELINE.json's north_star states:
loss within 1e-3 relative, dX / dW cosine similarity >= 0.999, argmax bit-exact."""
import numpy as np
import pytest
import torch

from conftest import cosine_similarity
from ddp_classification_pytorch_b200.head import CudaEngine  # noqa: E402
from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib
from oracle import arcface_oracle as orc

pytestmark = pytest.mark.gpu

LOSS_RTOL = 1e-3
GRAD_COS = 0.999


def _module(w, s, m, easy, **kw):
    C, D = w.shape
    head = ArcMarginProduct(D, C, s=s, m=m, easy_margin=easy, **kw).cuda()
    with torch.no_grad():
        head.weight.copy_(torch.as_tensor(w))
    return head


def _sync_ok():
    _lib.check(_lib.lib().arcb_check_watchdog(torch.cuda.current_stream().cuda_stream))


def test_fused_loss_matches_reference_golden(golden):
    g = golden
    head = _module(g["w"], float(g["s"]), float(g["m"]), bool(g["easy_margin"]))
    x = torch.tensor(g["x"], device="cuda", requires_grad=True)
    y = torch.tensor(g["label"], device="cuda")
    loss = head.forward_loss(x, y)
    loss.backward()
    _sync_ok()
    assert loss.item() == pytest.approx(float(g["loss"]), rel=LOSS_RTOL)
    assert cosine_similarity(x.grad.cpu().numpy(), g["dx"]) >= GRAD_COS
    assert cosine_similarity(head.weight.grad.cpu().numpy(), g["dw"]) >= GRAD_COS
    # magnitude as well as direction
    assert np.linalg.norm(x.grad.cpu().numpy()) == pytest.approx(np.linalg.norm(g["dx"]), rel=2e-2)
    assert np.linalg.norm(head.weight.grad.cpu().numpy()) == pytest.approx(np.linalg.norm(g["dw"]), rel=2e-2)


def test_logits_forward_matches_reference_golden(golden):
    g = golden
    s = float(g["s"])
    head = _module(g["w"], s, float(g["m"]), bool(g["easy_margin"]))
    x = torch.tensor(g["x"], device="cuda", requires_grad=True)
    y = torch.tensor(g["label"], device="cuda")
    out = head(x, y)
    assert out.dtype == torch.float32 and tuple(out.shape) == g["logits"].shape
    _sync_ok()
    o = out.detach().cpu().numpy()
    rows = np.arange(o.shape[0])
    # target column comes from the exact fp32 side path
    np.testing.assert_allclose(o[rows, g["label"]], g["logits"][rows, g["label"]], atol=2e-4 * s / 30, rtol=1e-5)
    # other columns: bf16 operands, fp32 accumulate  ->  |dcos| <= ~2^-8 worst case
    np.testing.assert_allclose(o, g["logits"], atol=4e-3 * s)
    # argmax must be identical wherever the fp32 reference is unambiguous at bf16 resolution
    srt = np.sort(g["logits"], axis=1)
    clear = (srt[:, -1] - srt[:, -2]) > 8e-3 * s
    assert (o.argmax(1)[clear] == g["logits"].argmax(1)[clear]).all()
    # the materialising path is differentiable like the reference's
    loss = torch.nn.functional.cross_entropy(out, y)
    loss.backward()
    _sync_ok()
    assert loss.item() == pytest.approx(float(g["loss"]), rel=LOSS_RTOL)
    assert cosine_similarity(x.grad.cpu().numpy(), g["dx"]) >= GRAD_COS
    assert cosine_similarity(head.weight.grad.cpu().numpy(), g["dw"]) >= GRAD_COS


@pytest.mark.parametrize("easy", [False, True])
@pytest.mark.parametrize("shape", [(256, 512, 5000), (300, 256, 2173), (512, 2048, 1000), (64, 100, 700)])
def test_fused_vs_oracle_medium(shape, easy):
    B, D, C = shape
    gen = torch.Generator().manual_seed(B + C)
    bound = (6.0 / (C + D)) ** 0.5
    w = (torch.rand(C, D, generator=gen) * 2 - 1) * bound
    x = torch.randn(B, D, generator=gen)
    y = torch.randint(0, C, (B,), generator=gen)
    # plant rows on both sides of every margin branch
    for r in range(0, B, 7):
        wy = w[y[r]] / w[y[r]].norm()
        c = [-0.99, -0.9, -0.86, -0.01, 0.01, 0.5, 0.99][(r // 7) % 7]
        n = torch.randn(D, generator=gen); n = n - (n @ wy) * wy; n = n / n.norm()
        x[r] = 3.0 * (c * wy + (1 - c * c) ** 0.5 * n)
    ref = orc.head_forward_backward(x.numpy(), w.numpy(), y.numpy(), 30.0, 0.5, easy)
    head = _module(w.numpy(), 30.0, 0.5, easy, chunk_classes=1024)  # several backward chunks
    xg = x.cuda().requires_grad_(True)
    loss = head.forward_loss(xg, y.cuda())
    loss.backward()
    _sync_ok()
    assert loss.item() == pytest.approx(ref["loss"], rel=LOSS_RTOL)
    assert cosine_similarity(xg.grad.cpu().numpy(), ref["dx"]) >= GRAD_COS
    assert cosine_similarity(head.weight.grad.cpu().numpy(), ref["dw"]) >= GRAD_COS
    np.testing.assert_allclose(head.aux["lse"].cpu().numpy(), ref["lse"], rtol=2e-3)
    np.testing.assert_allclose(head.aux["target_logit"].cpu().numpy(), ref["target_logit"], atol=1e-3, rtol=1e-5)


def test_grad_scale_and_label_dtypes():
    g = np.load(__import__("os").path.join(__import__("conftest").GOLDEN_DIR, "small_easy.npz"))
    head = _module(g["w"], 30.0, 0.5, True)
    x = torch.tensor(g["x"], device="cuda", requires_grad=True)
    base = None
    for lab in (torch.tensor(g["label"]).int(), torch.tensor(g["label"]).view(-1, 1), torch.tensor(g["label"]).float()):
        head.weight.grad = None; x.grad = None
        loss = head.forward_loss(x, lab.cuda())
        (loss * -2.5).backward()
        _sync_ok()
        if base is None:
            base = (loss.item(), x.grad.clone(), head.weight.grad.clone())
            assert cosine_similarity(x.grad.cpu().numpy(), -g["dx"]) >= GRAD_COS
            assert np.linalg.norm(x.grad.cpu().numpy()) == pytest.approx(2.5 * np.linalg.norm(g["dx"]), rel=2e-2)
        else:
            assert loss.item() == base[0]
            # split-K / projection sums use fp32 atomics, so the summation order may differ run to run
            torch.testing.assert_close(x.grad, base[1], rtol=1e-4, atol=1e-6)
            torch.testing.assert_close(head.weight.grad, base[2], rtol=1e-4, atol=1e-6)


def test_contract_errors():
    head = ArcMarginProduct(16, 8).cuda()
    with pytest.raises(IndexError):
        head(torch.randn(16, device="cuda"), torch.zeros(1, device="cuda"))
    with pytest.raises(RuntimeError):
        head(torch.randn(4, 16), torch.zeros(4))
    assert list(head.state_dict().keys()) == ["weight"]
    assert head.weight.dtype == torch.float32 and tuple(head.weight.shape) == (8, 16)


@pytest.mark.parametrize("chunk", [128, 384, 4096])
def test_persistent_and_per_chunk_backward_agree(chunk):
    """The persistent cooperative backward and the per-chunk launch sequence are two schedules of the
    same arithmetic: gradients must agree to fp32-atomic-order noise, for 1, few and many chunks."""
    B, D, C = 300, 192, 3000
    gen = torch.Generator().manual_seed(7)
    w = (torch.rand(C, D, generator=gen) * 2 - 1) * 0.05
    x = torch.randn(B, D, generator=gen)
    y = torch.randint(0, C, (B,), generator=gen)
    grads = []
    # claimed tiles (default), one launch per GEMM and chunk, fixed share of the tile list, one CTA per tile:
    # per-head flags, so the four heads coexist in one process
    for flags in (0, _lib.BWD_RECOMPUTE, _lib.BWD_PER_CHUNK, _lib.BWD_STATIC_SHARE, _lib.BWD_SINGLE_CTA):
        head = _module(w.numpy(), 30.0, 0.5, False, chunk_classes=chunk, kernel_flags=flags)
        xg = x.cuda().requires_grad_(True)
        head.forward_loss(xg, y.cuda()).backward()
        _sync_ok()
        grads.append((xg.grad.clone(), head.weight.grad.clone()))
    for other in grads[1:]:
        torch.testing.assert_close(grads[0][0], other[0], rtol=1e-3, atol=1e-6)
        torch.testing.assert_close(grads[0][1], other[1], rtol=1e-3, atol=1e-6)
    ref = orc.head_forward_backward(x.numpy(), w.numpy(), y.numpy(), 30.0, 0.5, False)
    assert cosine_similarity(grads[0][0].cpu().numpy(), ref["dx"]) >= GRAD_COS
    assert cosine_similarity(grads[0][1].cpu().numpy(), ref["dw"]) >= GRAD_COS


@pytest.mark.parametrize("shape", [(300, 192, 3000), (64, 100, 517), (520, 512, 9000), (128, 2048, 700)])
def test_weight_normalised_inside_the_gemm_matches_separate_kernel(shape):
    """F.normalize(weight) as extra warps of the forward GEMM kernel vs its own kernel: the same w_hat / inv_norm
    bits, the same statistics (the target cosine takes its row norm from another reduction order: 1e-6), the same
    gradients -- for ragged sizes, several tiles per CTA and every register-path width."""
    import ctypes
    B, D, C = shape
    gen = torch.Generator().manual_seed(11)
    w = ((torch.rand(C, D, generator=gen) * 2 - 1) * 0.05).cuda()
    w[3] = 0  # a clamped row (norm below eps)
    x = torch.randn(B, D, generator=gen).cuda()
    y = torch.randint(0, C, (B,), generator=gen).cuda()
    L = _lib.lib()
    eng = CudaEngine()
    Dp = L.arcb_padded_dim(D)
    out = []
    for flags in (_lib.FWD_NORM_FUSED, _lib.FWD_NORM_SEPARATE):
        cfg = {"s": 30.0, "m": 0.5, "easy_margin": False, "class_offset": 0, "chunk_classes": 512, "flags": flags}
        n0 = L.arcb_launch_count()
        stats, state = eng.shard_forward(x, w, y, cfg)
        launches = L.arcb_launch_count() - n0
        lse, zt, loss = eng.merge_loss(stats, 1, B)
        dw, dxhat = eng.shard_backward(state, cfg, lse=lse, zt=zt)
        _sync_ok()
        ws = state[3]
        a, b = ctypes.c_void_p(), ctypes.c_void_p()
        _lib.check(L.arcb_ws_weight_slots(ws.data_ptr(), C, D, ctypes.byref(a), ctypes.byref(b)))
        off_a, off_b = a.value - ws.data_ptr(), b.value - ws.data_ptr()
        what = ws[off_a: off_a + C * Dp * 2].view(torch.bfloat16).view(C, Dp).clone()
        inv = ws[off_b: off_b + C * 4].view(torch.float32).clone()
        out.append((stats.clone(), what, inv, loss.clone(), dw, dxhat, launches))
    f, u = out
    assert f[6] == u[6] - 1                      # the normalisation launch is gone
    assert torch.equal(f[1].view(torch.int16), u[1].view(torch.int16))
    assert torch.equal(f[2], u[2])
    torch.testing.assert_close(f[0], u[0], rtol=2e-6, atol=1e-6)
    torch.testing.assert_close(f[3], u[3], rtol=1e-6, atol=0)
    torch.testing.assert_close(f[4], u[4], rtol=1e-3, atol=1e-7)
    torch.testing.assert_close(f[5], u[5], rtol=1e-3, atol=1e-7)


def _random_shapes(n, seed):
    rng = np.random.default_rng(seed)
    out = []
    for _ in range(n):
        B = int(rng.integers(1, 700))
        D = int(rng.choice([8, 36, 64, 100, 128, 200, 256, 384, 512]))
        C = int(rng.integers(1, 6000))
        chunk = int(rng.choice([128, 256, 512, 1024, 2048]))
        out.append((B, D, C, chunk, bool(rng.integers(0, 2)), float(rng.choice([16.0, 30.0, 64.0]))))
    return out


_EDGE_SHAPES = [(1, 64, 1, 128, False, 30.0), (3, 70, 5, 128, True, 30.0), (129, 66, 257, 256, False, 30.0),
                (2, 512, 2, 2048, False, 64.0)]


@pytest.mark.parametrize("case", _random_shapes(10, seed=2024) + _EDGE_SHAPES)
def test_random_shapes_vs_oracle(case):
    """Seeded sweep over ragged batch / feature / class counts and chunk lengths (tiles that are partly or wholly
    padding, shards smaller than one tile, many short chunks for the tile list and the claiming logic)."""
    B, D, C, chunk, easy, s_ = case
    gen = torch.Generator().manual_seed(B * 7919 + C)
    w = (torch.rand(C, D, generator=gen) * 2 - 1) * 0.1
    x = torch.randn(B, D, generator=gen)
    y = torch.randint(0, C, (B,), generator=gen)
    head = _module(w.numpy(), s_, 0.5, easy, chunk_classes=chunk)
    xg = x.cuda().requires_grad_(True)
    loss = head.forward_loss(xg, y.cuda())
    loss.backward()
    _sync_ok()
    ref = orc.head_forward_backward(x.numpy(), w.numpy(), y.numpy(), s_, 0.5, easy)
    assert abs(loss.item() - ref["loss"]) <= LOSS_RTOL * abs(ref["loss"])
    if np.linalg.norm(ref["dx"]) > 0:
        assert cosine_similarity(xg.grad.cpu().numpy(), ref["dx"]) >= GRAD_COS
        assert cosine_similarity(head.weight.grad.cpu().numpy(), ref["dw"]) >= GRAD_COS
        assert np.linalg.norm(head.weight.grad.cpu().numpy()) == pytest.approx(np.linalg.norm(ref["dw"]), rel=3e-2)


def _check_top3(head_aux, ref_logits, s):
    """argmax / top-3 must be IDENTICAL to the fp32 reference wherever the reference itself is unambiguous
    (neighbouring logits further apart than fp32 summation-order noise)."""
    ref_logits = np.asarray(ref_logits, dtype=np.float64)
    order = np.argsort(-ref_logits, axis=1, kind="stable")
    k = min(4, ref_logits.shape[1])
    top = np.take_along_axis(ref_logits, order[:, :k], axis=1)
    gaps = top[:, :-1] - top[:, 1:]
    clear = (gaps > 2e-5 * s).all(axis=1)
    assert clear.mean() > 0.9
    got = head_aux["top3"].cpu().numpy()
    want = order[:, : got.shape[1]]
    n = min(3, ref_logits.shape[1])
    assert (got[clear][:, :n] == want[clear][:, :n]).all()
    assert (head_aux["argmax"].cpu().numpy()[clear] == want[clear][:, 0]).all()
    np.testing.assert_allclose(head_aux["top3_logits"].cpu().numpy()[clear][:, :n], top[clear][:, :n], atol=2e-4 * s / 30,
                               rtol=1e-5)


def test_exact_top3_matches_reference_golden(golden):
    g = golden
    s = float(g["s"])
    head = _module(g["w"], s, float(g["m"]), bool(g["easy_margin"]))
    x = torch.tensor(g["x"], device="cuda")
    y = torch.tensor(g["label"], device="cuda")
    with torch.no_grad():
        head.forward_loss(x, y, want_top3=True)
    _sync_ok()
    _check_top3(head.aux, g["logits"], s)


@pytest.mark.parametrize("shape", [(512, 512, 40000), (200, 96, 3000)])
def test_exact_top3_vs_oracle_many_classes(shape):
    # at 40k classes ~10 % of the rows have a top-1/top-2 gap below bf16 resolution: the re-score must fix them
    B, D, C = shape
    gen = torch.Generator().manual_seed(C)
    w = (torch.rand(C, D, generator=gen) * 2 - 1) * (6.0 / (C + D)) ** 0.5
    x = torch.randn(B, D, generator=gen)
    y = torch.randint(0, C, (B,), generator=gen)
    ref = orc.head_forward(x.numpy(), w.numpy(), y.numpy(), 30.0, 0.5, True)
    head = _module(w.numpy(), 30.0, 0.5, True)
    with torch.no_grad():
        loss = head.forward_loss(x.cuda(), y.cuda(), want_top3=True)
    _sync_ok()
    assert torch.isfinite(loss)
    _check_top3(head.aux, ref, 30.0)


@pytest.mark.parametrize("shape", [(1000, 512, 20000), (520, 192, 40003), (300, 64, 80000)])
@pytest.mark.parametrize("fused_norm", [False, True])
def test_resident_forward_matches_streaming_forward(shape, fused_norm):
    """The x_hat-resident forward schedule (sample block kept in shared memory, one partial per row and CTA) against
    the streaming one (a partial per row and class block + merge): same statistics, same loss, same exact top-3, and
    both equal the oracle -- ragged class counts, ragged batch, with and without the in-kernel weight normalisation."""
    B, D, C = shape
    gen = torch.Generator().manual_seed(B + C)
    w = ((torch.rand(C, D, generator=gen) * 2 - 1) * 0.05).cuda()
    x = torch.randn(B, D, generator=gen).cuda()
    y = torch.randint(0, C, (B,), generator=gen).cuda()
    x[:8] = w[y[:8]] * 50 + 0.05 * x[:8]          # rows next to their class: large target logits, both branches
    L = _lib.lib()
    eng = CudaEngine()
    norm = _lib.FWD_NORM_FUSED if fused_norm else _lib.FWD_NORM_SEPARATE
    res = {}
    norm |= _lib.BWD_RECOMPUTE  # (the stored-probabilities default has its own forward epilogue on the streaming core)
    for name, flags in (("resident", norm), ("streaming", norm | _lib.FWD_STREAMING)):
        cfg = {"s": 30.0, "m": 0.5, "easy_margin": False, "class_offset": 0, "chunk_classes": 0, "flags": flags,
               "want_top3": True}
        n0 = L.arcb_launch_count()
        stats, state = eng.shard_forward(x, w, y, cfg)
        launches = L.arcb_launch_count() - n0
        lse, zt, loss = eng.merge_loss(stats, 1, B)
        val, idx = eng.shard_top3(x, w, state, cfg)
        dw, dxhat = eng.shard_backward(state, cfg, lse=lse, zt=zt)
        _sync_ok()
        res[name] = (lse.clone(), zt.clone(), loss.item(), val.clone(), idx.clone(), dw, dxhat, launches)
    r, s_ = res["resident"], res["streaming"]
    assert r[7] <= s_[7]                                         # at most one merge launch (two-level at > 32k classes)
    torch.testing.assert_close(r[0], s_[0], rtol=1e-5, atol=1e-5)  # lse: same bf16 cosines, other summation order
    assert torch.equal(r[1], s_[1])                                # target logit: the same fp32 side path
    assert r[2] == pytest.approx(s_[2], rel=1e-5)
    assert torch.equal(r[4], s_[4])                                # exact top-3 from either set of block maxima
    torch.testing.assert_close(r[3], s_[3], rtol=1e-6, atol=1e-6)
    torch.testing.assert_close(r[5], s_[5], rtol=1e-3, atol=1e-6)   # two backward runs: fp32 atomic order on tiny entries
    ref = orc.head_forward_backward(x.cpu().numpy(), w.cpu().numpy(), y.cpu().numpy(), 30.0, 0.5, False)
    assert abs(r[2] - ref["loss"]) <= LOSS_RTOL * abs(ref["loss"])
    np.testing.assert_allclose(r[0].cpu().numpy(), ref["lse"], rtol=2e-3)
    assert cosine_similarity(r[5].cpu().numpy(), ref["dw"]) >= GRAD_COS


@pytest.mark.parametrize("shape,easy", [((1000, 512, 20000), False), ((520, 320, 40003), True), ((1024, 512, 150000), False),
                                        ((340, 318, 50000), False)])  # D % 4 != 0: dW leaves by plain stores (EpiDWT<false>)
def test_backward_from_stored_probabilities_matches_recompute_and_oracle(shape, easy):
    """Stored probabilities (the default for these shapes): the forward keeps bf16 softmax numerators and the backward runs
    as two plain GEMMs (no recompute GEMM); ARCB_BWD_RECOMPUTE selects the recompute schedule.  Same loss, gradients within the parity bar of the oracle and of the default recompute path, planted rows on
    both sides of the margin branches, ragged class counts and feature sizes."""
    B, D, C = shape
    gen = torch.Generator().manual_seed(B * 7 + C)
    w = ((torch.rand(C, D, generator=gen) * 2 - 1) * 0.05)
    x = torch.randn(B, D, generator=gen)
    y = torch.randint(0, C, (B,), generator=gen)
    for r, c in zip(range(0, 64, 4), [-0.99, -0.9, -0.86, -0.3, -0.01, 0.01, 0.4, 0.9] * 2):  # target cosines to plant
        wy = w[y[r]] / w[y[r]].norm()
        n = torch.randn(D, generator=gen); n = n - (n @ wy) * wy; n = n / n.norm()
        x[r] = 2.0 * (c * wy + (1 - c * c) ** 0.5 * n)
    L = _lib.lib()
    assert L.arcb_keep_probs_active(B, C, D, 30.0, 0) == 1
    assert L.arcb_keep_probs_active(B, C, D, 30.0, _lib.BWD_RECOMPUTE) == 0 and L.arcb_keep_probs_active(B, C, D, 64.0, 0) == 0
    assert L.arcb_keep_probs_active(B, C, D, 30.0, _lib.BWD_PER_CHUNK) == 0 and L.arcb_keep_probs_active(B, C, 256, 30.0, 0) == 0
    out = {}
    for name, flags in (("probs", 0), ("recompute", _lib.BWD_RECOMPUTE)):
        head = _module(w.numpy(), 30.0, 0.5, easy, kernel_flags=flags)
        xg = x.cuda().requires_grad_(True)
        loss = head.forward_loss(xg, y.cuda(), want_top3=(name == "probs"))
        (loss * 1.5).backward()
        _sync_ok()
        out[name] = (loss.item(), xg.grad.clone(), head.weight.grad.clone(), head.aux["lse"].clone())
    p, r = out["probs"], out["recompute"]
    assert p[0] == pytest.approx(r[0], rel=1e-4)
    torch.testing.assert_close(p[3], r[3], rtol=1e-4, atol=1e-4)
    assert cosine_similarity(p[1].cpu().numpy(), r[1].cpu().numpy()) >= 0.9995
    assert cosine_similarity(p[2].cpu().numpy(), r[2].cpu().numpy()) >= 0.9995
    if C <= 50000:  # the fp64 oracle finishes in seconds up to here
        ref = orc.head_forward_backward(x.numpy(), w.numpy(), y.numpy(), 30.0, 0.5, easy, grad_scale=1.5)
        assert abs(p[0] - ref["loss"]) <= LOSS_RTOL * abs(ref["loss"])
        assert cosine_similarity(p[1].cpu().numpy(), ref["dx"]) >= GRAD_COS
        assert cosine_similarity(p[2].cpu().numpy(), ref["dw"]) >= GRAD_COS
        assert np.linalg.norm(p[2].cpu().numpy()) == pytest.approx(np.linalg.norm(ref["dw"]), rel=2e-2)
        assert np.linalg.norm(p[1].cpu().numpy()) == pytest.approx(np.linalg.norm(ref["dx"]), rel=2e-2)
        # F.normalize backward: every gradient row is orthogonal to its own input row (the t_j swap got the projection right)
        wn = torch.nn.functional.normalize(w.cuda().double(), dim=1)
        gw = p[2].double()
        rel = (wn * gw).sum(1).abs() / (gw.norm(dim=1) + 1e-30)
        assert rel.median().item() < 2e-2
