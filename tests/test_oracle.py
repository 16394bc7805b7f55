"""The oracle against the reference's own outputs (committed golden vectors) and KATs.

CPU only.  The golden vectors were produced by running the reference class itself
(tests/golden/make_golden.py); when /root/reference is present the live reference is
checked too.
"""
import numpy as np
import pytest

from conftest import cosine_similarity, load_golden
from oracle import arcface_oracle as orc
from oracle import ref_extract


def test_constants_m05():
    k = orc.margin_constants(30.0, 0.5, False)  # arc_main.py:149-152
    assert k.cos_m == pytest.approx(0.8775825619, abs=1e-9)
    assert k.sin_m == pytest.approx(0.4794255386, abs=1e-9)
    assert k.th == pytest.approx(-0.8775825619, abs=1e-9)
    assert k.mm == pytest.approx(0.2397127693, abs=1e-9)


def test_oracle_matches_reference_golden(golden):
    g = golden
    out = orc.head_forward_backward(g["x"], g["w"], g["label"], float(g["s"]), float(g["m"]), bool(g["easy_margin"]))
    # fp64 restatement vs fp32 reference: differences are fp32 rounding of the reference
    np.testing.assert_allclose(out["logits"], g["logits"], rtol=0, atol=3e-5 * float(g["s"]) / 30.0 * 2)
    assert out["loss"] == pytest.approx(float(g["loss"]), rel=2e-6)
    assert cosine_similarity(out["dx"], g["dx"]) > 1 - 1e-9
    assert cosine_similarity(out["dw"], g["dw"]) > 1 - 1e-9
    np.testing.assert_allclose(out["dx"], g["dx"], rtol=0, atol=2e-5 * np.abs(g["dx"]).max() + 1e-9)
    np.testing.assert_allclose(out["dw"], g["dw"], rtol=0, atol=2e-5 * np.abs(g["dw"]).max() + 1e-9)
    # argmax / target index are integer results: must be identical wherever the fp32 reference
    # itself is unambiguous (top-2 gap above fp32 noise)
    ref_sorted = np.sort(g["logits"], axis=1)
    clear = (ref_sorted[:, -1] - ref_sorted[:, -2]) > 1e-4
    assert (out["argmax"][clear] == g["logits"].argmax(1)[clear]).all()


def test_known_answers():
    k = load_golden("kat")
    for easy, key in ((False, "logits_hard"), (True, "logits_easy")):
        z = orc.head_forward(k["x"], k["w"], k["label"], 30.0, 0.5, easy)
        np.testing.assert_allclose(z, k[key], atol=2e-5)
    z = orc.head_forward(k["x"], k["w"], k["label"], 30.0, 0.5, False) / 30.0
    assert z[0, 0] == pytest.approx(0.8775825619)      # cos_t=+1 -> cos_m
    assert z[1, 1] == pytest.approx(-1.2397127693)     # cos_t=-1 hard -> -1-mm
    assert z[2, 2] == pytest.approx(-0.4794255386)     # zero row: cos=0 > th -> -sin_m
    assert z[1, 0] == pytest.approx(0.0) and z[0, 1] == pytest.approx(0.0)
    ze = orc.head_forward(k["x"], k["w"], k["label"], 30.0, 0.5, True) / 30.0
    assert ze[1, 1] == pytest.approx(-1.0) and ze[2, 2] == pytest.approx(0.0)


def test_label_contract():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((6, 16)); w = rng.standard_normal((5, 16)); y = rng.integers(0, 5, 6)
    base = orc.head_forward(x, w, y)
    for lab in (y.astype(np.int32), y.reshape(-1, 1), y.astype(np.float32)):  # arc_main.py:169
        np.testing.assert_array_equal(orc.head_forward(x, w, lab), base)
    with pytest.raises(IndexError):
        orc.head_forward(x, w, np.full(6, 5))
    with pytest.raises(IndexError):
        orc.head_forward(x[0], w, y)


def test_finite_grads_at_unit_cosine():
    # the reference yields NaN here (sqrt'(0) * 0); ours is defined to stay finite
    w = np.eye(3, 8); x = np.stack([-2.0 * w[1], w[0]]); y = np.array([1, 0])
    out = orc.head_forward_backward(x, w, y)
    assert np.isfinite(out["dx"]).all() and np.isfinite(out["dw"]).all()


def test_backward_against_numeric_gradient():
    rng = np.random.default_rng(5)
    x = rng.standard_normal((4, 6)); w = rng.standard_normal((7, 6)) * 0.3; y = np.array([1, 6, 0, 3])
    for easy in (False, True):
        out = orc.head_forward_backward(x, w, y, 8.0, 0.4, easy)
        eps = 1e-6
        for arr, grad in ((x, out["dx"]), (w, out["dw"])):
            for idx in [(0, 0), (1, 3), (3, 5)]:
                old = arr[idx]
                arr[idx] = old + eps; lp = orc.head_forward_backward(x, w, y, 8.0, 0.4, easy)["loss"]
                arr[idx] = old - eps; lm = orc.head_forward_backward(x, w, y, 8.0, 0.4, easy)["loss"]
                arr[idx] = old
                assert grad[idx] == pytest.approx((lp - lm) / (2 * eps), rel=1e-4, abs=1e-7)


def test_shard_merge_equals_unsharded():
    rng = np.random.default_rng(2)
    B, D, C = 16, 32, 50
    x = rng.standard_normal((B, D)); w = rng.standard_normal((C, D)); y = rng.integers(0, C, B)
    full = orc.head_forward_backward(x, w, y, 30.0, 0.5, True)
    for parts in (2, 3, 5):
        bounds = np.linspace(0, C, parts + 1).astype(int)
        recs = [orc.shard_forward_stats(x, w[a:b], y, a, 30.0, 0.5, True) for a, b in zip(bounds[:-1], bounds[1:])]
        lse, tgt, loss = orc.merge_shard_stats(*[np.stack(r) for r in zip(*recs)])
        np.testing.assert_allclose(lse, full["lse"], rtol=1e-12)
        np.testing.assert_allclose(tgt, full["target_logit"], rtol=1e-12)
        assert loss == pytest.approx(full["loss"], rel=1e-12)


def test_torch_port_matches_oracle():
    import torch
    g = load_golden("small_easy")
    port = orc.TorchPort(torch.from_numpy(g["w"]), float(g["s"]), float(g["m"]), bool(g["easy_margin"]))
    loss, dx, dw = port.step(torch.from_numpy(g["x"]), torch.from_numpy(g["label"]))
    assert loss == pytest.approx(float(g["loss"]), rel=1e-6)
    np.testing.assert_allclose(dx.numpy(), g["dx"], atol=1e-6 * np.abs(g["dx"]).max() + 1e-9)
    np.testing.assert_allclose(dw.numpy(), g["dw"], atol=1e-6 * np.abs(g["dw"]).max() + 1e-9)


@pytest.mark.skipif(not ref_extract.reference_available(), reason="/root/reference only exists in the build container")
def test_oracle_matches_live_reference():
    import torch
    torch.manual_seed(11)
    x = torch.randn(24, 48); w = torch.randn(77, 48) * 0.2; y = torch.randint(0, 77, (24,))
    for easy in (False, True):
        ref = ref_extract.reference_forward_backward(x, w, y, 30.0, 0.5, easy)
        out = orc.head_forward_backward(x.numpy(), w.numpy(), y.numpy(), 30.0, 0.5, easy)
        assert out["loss"] == pytest.approx(ref["loss"], rel=2e-6)
        np.testing.assert_allclose(out["logits"], ref["logits"].numpy(), atol=6e-5)
        assert cosine_similarity(out["dx"], ref["dx"].numpy()) > 1 - 1e-9
        assert cosine_similarity(out["dw"], ref["dw"].numpy()) > 1 - 1e-9
