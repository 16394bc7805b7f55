"""GPU parity of the fused optimizer step (arcb_sgd_step / arcb_adam_step) and of the cached-normalisation forward.

Checked against (1) the committed torch.optim fixtures, (2) oracle/optim_oracle.py, (3) torch.optim on the GPU,
and (4) the un-cached head, all through the C ABI / the public module.
"""
import ast
import glob
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import optim_oracle as oo  # noqa: E402

OPT_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "optim")
CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(OPT_DIR, "*.npz")))


def load_case(name):
    with np.load(os.path.join(OPT_DIR, name + ".npz")) as z:
        d = {k: z[k] for k in z.files}
    d["kind"] = str(d["kind"])
    d["kwargs"] = ast.literal_eval(str(d["kwargs"]))
    return d


def _raw_step(kind, kw, w, g, s0, s1, step, what=None, inv=None):
    """One call through the C ABI on CUDA tensors (in place)."""
    from ddp_classification_pytorch_b200 import _lib
    L = _lib.lib()
    rows, D = w.shape
    Dp = L.arcb_padded_dim(D)
    st = torch.cuda.current_stream().cuda_stream
    p = lambda t: None if t is None else t.data_ptr()  # noqa: E731
    if kind == "sgd":
        _lib.check(L.arcb_sgd_step(p(w), D, p(g), D, p(s0), D, rows, D, kw["lr"], kw.get("momentum", 0.0),
                                   kw.get("dampening", 0.0), kw.get("weight_decay", 0.0),
                                   int(kw.get("nesterov", False)), int(step == 1), p(what), Dp, p(inv), st))
    else:
        b1, b2 = kw.get("betas", (0.9, 0.999))
        _lib.check(L.arcb_adam_step(p(w), D, p(g), D, p(s0), p(s1), D, rows, D, kw["lr"], b1, b2, kw.get("eps", 1e-8),
                                    kw.get("weight_decay", 0.0), step, p(what), Dp, p(inv), st))


@pytest.mark.parametrize("name", CASES)
def test_fixture_trajectory(name):
    """Whole trajectory on the GPU vs torch.optim's recorded one; per-step vs the fp64 oracle; w_hat vs oracle."""
    from ddp_classification_pytorch_b200 import _lib
    d = load_case(name)
    kind, kw = d["kind"], d["kwargs"]
    dev = "cuda"
    w = torch.tensor(d["w0"], device=dev)
    rows, D = w.shape
    Dp = _lib.lib().arcb_padded_dim(D)
    s0 = torch.zeros_like(w) if (kind == "adam" or kw.get("momentum", 0.0) != 0.0) else None
    s1 = torch.zeros_like(w) if kind == "adam" else None
    what = torch.full((rows, Dp), float("nan"), device=dev, dtype=torch.bfloat16)
    inv = torch.empty(rows, device=dev)
    for t in range(d["grads"].shape[0]):
        g = torch.tensor(d["grads"][t], device=dev)
        w_before = w.cpu().numpy()
        s0_before = None if s0 is None else s0.cpu().numpy()
        s1_before = None if s1 is None else s1.cpu().numpy()
        _raw_step(kind, kw, w, g, s0, s1, t + 1, what, inv)
        torch.cuda.synchronize()
        # (1) torch.optim fixture (fp32 both sides: a few ulp from fma / ordering differences)
        np.testing.assert_allclose(w.cpu().numpy(), d["w"][t], rtol=3e-6, atol=3e-8)
        if s0 is not None:
            np.testing.assert_allclose(s0.cpu().numpy(), d["s0"][t], rtol=3e-6, atol=5e-8)
        if s1 is not None:
            np.testing.assert_allclose(s1.cpu().numpy(), d["s1"][t], rtol=3e-6, atol=1e-11)
        # (2) fp64 oracle, one step from the kernel's own previous state
        if kind == "sgd":
            ow, _ = oo.sgd_step(w_before, d["grads"][t], None if t == 0 else s0_before, **kw)
        else:
            ow, _, _ = oo.adam_step(w_before, d["grads"][t], s0_before, s1_before, t + 1, **kw)
        np.testing.assert_allclose(w.cpu().numpy(), ow, rtol=2e-6, atol=2e-8)
        # (3) the normalised copy of the UPDATED weight
        oh, oinv = oo.normalize_rows(w.cpu().numpy().astype(np.float64))
        np.testing.assert_allclose(what[:, :D].float().cpu().numpy(), oh, rtol=0, atol=2 ** -8)  # bf16 rounding
        assert torch.all(what[:, D:].float() == 0)
        np.testing.assert_allclose(inv.cpu().numpy(), oinv, rtol=2e-6)


@pytest.mark.parametrize("D", [64, 100, 256, 512, 1024, 2048, 2500])
@pytest.mark.parametrize("kind", ["sgd", "adam"])
def test_against_torch_optim_on_gpu(kind, D):
    """Every register-path width (and the generic path) vs torch.optim on the same device, 3 steps."""
    from ddp_classification_pytorch_b200 import _lib
    torch.manual_seed(D)
    rows = 301
    kw = dict(lr=0.01, momentum=0.9, weight_decay=5e-4) if kind == "sgd" else dict(lr=1e-3, weight_decay=1e-4)
    w0 = torch.randn(rows, D, device="cuda") * 0.05
    ref = torch.nn.Parameter(w0.clone())
    opt = (torch.optim.SGD if kind == "sgd" else torch.optim.Adam)([ref], **kw)
    w = w0.clone()
    s0, s1 = torch.zeros_like(w), (torch.zeros_like(w) if kind == "adam" else None)
    Dp = _lib.lib().arcb_padded_dim(D)
    what = torch.empty(rows, Dp, device="cuda", dtype=torch.bfloat16)
    inv = torch.empty(rows, device="cuda")
    what2 = torch.empty_like(what)
    inv2 = torch.empty_like(inv)
    for t in range(3):
        g = torch.randn(rows, D, device="cuda") * 0.02
        ref.grad = g.clone()
        opt.step()
        _raw_step(kind, kw, w, g, s0, s1, t + 1, what, inv)
        torch.testing.assert_close(w, ref.detach(), rtol=3e-6, atol=3e-8)
    # the deposited copy equals what the stand-alone normalise kernel produces from the updated weight
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(_lib.lib().arcb_l2norm_rows(w.data_ptr(), rows, D, D, what2.data_ptr(), Dp, inv2.data_ptr(), st))
    torch.testing.assert_close(inv, inv2, rtol=1e-6, atol=0)
    assert (what.float() - what2.float()).abs().max().item() <= 2 ** -8


def test_without_normalised_output_and_plain_sgd():
    """w_hat = NULL: a plain fused optimizer; momentum == 0: no state buffer needed."""
    torch.manual_seed(0)
    w = torch.randn(50, 128, device="cuda")
    g = torch.randn_like(w)
    expect = w - 0.1 * g
    _raw_step("sgd", dict(lr=0.1), w, g, None, None, 1)
    torch.testing.assert_close(w, expect, rtol=1e-6, atol=1e-7)


def test_argument_errors():
    from ddp_classification_pytorch_b200 import _lib
    L = _lib.lib()
    w = torch.zeros(4, 64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    assert L.arcb_sgd_step(w.data_ptr(), 64, w.data_ptr(), 64, None, 64, 4, 64, 0.1, 0.9, 0.0, 0.0, 0, 1, None, 64, None,
                           st) == -1  # momentum without a buffer
    assert b"momentum" in L.arcb_last_error()
    assert L.arcb_adam_step(w.data_ptr(), 64, w.data_ptr(), 64, w.data_ptr(), w.data_ptr(), 64, 4, 64, 1e-3, 0.9, 0.999,
                            1e-8, 0.0, 0, None, 64, None, st) == -1  # step counts from 1
    assert L.arcb_sgd_step(w.data_ptr(), 64, w.data_ptr(), 64, None, 64, 4, 64, 0.1, 0.0, 0.0, 0.0, 0, 1, w.data_ptr(),
                           64, None, st) == -1  # w_hat without inv_norm


def _make_head(C, D, **kw):
    from ddp_classification_pytorch_b200 import ArcMarginProduct
    torch.manual_seed(5)
    return ArcMarginProduct(D, C, s=30.0, m=0.5, easy_margin=False, **kw).cuda()


@pytest.mark.parametrize("kind", ["sgd", "adam"])
def test_training_loop_matches_torch_optim(kind):
    """HeadSGD / HeadAdam + cached normalisation vs the same head driven by torch.optim, 5 steps."""
    from ddp_classification_pytorch_b200 import HeadAdam, HeadSGD, _lib
    C, D, B = 3000, 128, 64
    a = _make_head(C, D)
    b = _make_head(C, D)
    b.load_state_dict(a.state_dict())
    if kind == "sgd":
        kw = dict(lr=0.05, momentum=0.9, weight_decay=5e-4)
        oa, ob = HeadSGD(a, **kw), torch.optim.SGD(b.parameters(), **kw)
    else:
        kw = dict(lr=1e-3)
        oa, ob = HeadAdam(a, **kw), torch.optim.Adam(b.parameters(), **kw)
    sched_a = torch.optim.lr_scheduler.StepLR(oa, step_size=2, gamma=0.1)  # arc_main.py:255
    sched_b = torch.optim.lr_scheduler.StepLR(ob, step_size=2, gamma=0.1)
    g = torch.Generator(device="cuda").manual_seed(1)
    L = _lib.lib()
    for it in range(5):
        x = torch.randn(B, D, device="cuda", generator=g)
        y = torch.randint(0, C, (B,), device="cuda", generator=g)
        for head, opt in ((a, oa), (b, ob)):
            opt.zero_grad()
            loss = head.forward_loss(x, y)
            loss.backward()
            opt.step()
            head.last_loss = loss.item()
        sched_a.step()
        sched_b.step()
        assert abs(a.last_loss - b.last_loss) <= 1e-5 * abs(b.last_loss)
        torch.testing.assert_close(a.weight, b.weight, rtol=1e-5, atol=1e-7)
    # the cache was valid at every forward after the first: the deposit is what the forward consumed
    assert a._wcache.valid == (a.weight.data_ptr(), a.weight._version)
    assert L.arcb_check_watchdog(torch.cuda.current_stream().cuda_stream) == 0
    # state_dict interchange with the torch optimizer it replaces
    sd = oa.state_dict()
    ob.load_state_dict(sd)
    oa.load_state_dict(ob.state_dict())


def test_cached_forward_skips_normalise_and_stays_exact():
    """Launch accounting + bit-identical loss: cached (prenorm) forward vs the normalising forward."""
    from ddp_classification_pytorch_b200 import _lib
    C, D, B = 5000, 256, 128
    a = _make_head(C, D, cache_normalized_weight=True)
    b = _make_head(C, D)
    b.load_state_dict(a.state_dict())
    x = torch.randn(B, D, device="cuda")
    y = torch.randint(0, C, (B,), device="cuda")
    L = _lib.lib()
    with torch.no_grad():
        l_ref = b.forward_loss(x, y)
        l1 = a.forward_loss(x, y)          # fills the cache
        n0 = L.arcb_launch_count()
        l2 = a.forward_loss(x, y)          # must reuse it
        n1 = L.arcb_launch_count()
        lb = b.forward_loss(x, y)
        n2 = L.arcb_launch_count()
    assert (n2 - n1) - (n1 - n0) == 1      # exactly the weight normalisation launch is gone
    assert l1.item() == l_ref.item() == l2.item() == lb.item()
    # any in-place change of the weight invalidates the cache
    with torch.no_grad():
        a.weight.mul_(1.5)
        a.weight[7] += 0.3
        b.weight.copy_(a.weight)
        assert a.forward_loss(x, y).item() == b.forward_loss(x, y).item()
    # gradients through a cached workspace equal the un-cached ones
    xa = x.clone().requires_grad_(True)
    xb = x.clone().requires_grad_(True)
    a.forward_loss(xa, y).backward()
    b.forward_loss(xb, y).backward()
    torch.testing.assert_close(xa.grad, xb.grad, rtol=1e-5, atol=1e-8)
    torch.testing.assert_close(a.weight.grad, b.weight.grad, rtol=1e-5, atol=1e-9)


def test_pending_graph_keeps_its_workspace():
    """Two forwards before the first backward: the second must not overwrite the first one's saved state."""
    C, D, B = 2000, 64, 32
    a = _make_head(C, D, cache_normalized_weight=True)
    b = _make_head(C, D)
    b.load_state_dict(a.state_dict())
    x1 = torch.randn(B, D, device="cuda")
    x2 = torch.randn(B, D, device="cuda")
    y1 = torch.randint(0, C, (B,), device="cuda")
    y2 = torch.randint(0, C, (B,), device="cuda")
    la1 = a.forward_loss(x1, y1)
    assert a._wcache.busy == 1
    la2 = a.forward_loss(x2, y2)   # cache busy -> private workspace
    (la1 + la2).backward()
    assert a._wcache.busy == 0
    (b.forward_loss(x1, y1) + b.forward_loss(x2, y2)).backward()
    torch.testing.assert_close(a.weight.grad, b.weight.grad, rtol=1e-5, atol=1e-9)
    # a graph that is dropped without backward releases the workspace when it is collected
    l3 = a.forward_loss(x1, y1)
    assert a._wcache.busy == 1
    del l3
    assert a._wcache.busy == 0


def test_optimizer_rejects_cpu_and_wrapped_heads():
    from ddp_classification_pytorch_b200 import ArcMarginProduct, HeadSGD
    with pytest.raises(TypeError):
        HeadSGD(torch.nn.Linear(4, 4), lr=0.1)
    head = ArcMarginProduct(16, 10)
    opt = HeadSGD(head, lr=0.1)
    head.weight.grad = torch.zeros_like(head.weight)
    with pytest.raises(RuntimeError, match="CUDA"):
        opt.step()
