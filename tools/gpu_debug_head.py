"""Diagnostics for the fused head vs golden vectors / oracle: prints numbers, never asserts."""
import os, sys, glob, traceback
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib
from oracle import arcface_oracle as orc
from conftest import cosine_similarity as cs

def wd():
    rc = _lib.lib().arcb_check_watchdog(torch.cuda.current_stream().cuda_stream)
    return "" if rc == 0 else " WATCHDOG:" + _lib.lib().arcb_last_error().decode()

def run(name, x, w, y, s, m, easy, ref, chunk=None):
    C, D = w.shape
    try:
        head = ArcMarginProduct(D, C, s=s, m=m, easy_margin=easy, chunk_classes=chunk).cuda()
        with torch.no_grad(): head.weight.copy_(torch.as_tensor(w))
        xg = torch.tensor(x, device="cuda", requires_grad=True)
        yl = torch.tensor(y, device="cuda")
        loss = head.forward_loss(xg, yl); loss.backward(); msg = wd()
        dx = xg.grad.cpu().numpy(); dw = head.weight.grad.cpu().numpy()
        print(f"[fused] {name}: loss {loss.item():.6f} ref {ref['loss']:.6f} rel {abs(loss.item()-ref['loss'])/abs(ref['loss']):.2e} "
              f"cos(dx) {cs(dx, ref['dx']):.6f} cos(dw) {cs(dw, ref['dw']):.6f} "
              f"|dx| {np.linalg.norm(dx):.4e}/{np.linalg.norm(ref['dx']):.4e} |dw| {np.linalg.norm(dw):.4e}/{np.linalg.norm(ref['dw']):.4e}{msg}", flush=True)
        head.weight.grad = None; xg.grad = None
        out = head(xg, yl); l2 = torch.nn.functional.cross_entropy(out, yl); l2.backward(); msg = wd()
        o = out.detach().cpu().numpy()
        dx = xg.grad.cpu().numpy(); dw = head.weight.grad.cpu().numpy()
        extra = ""
        if "logits" in ref:
            extra = f" max|dlogit| {np.abs(o-ref['logits']).max():.3e}"
        print(f"[logits] {name}: loss {l2.item():.6f} cos(dx) {cs(dx, ref['dx']):.6f} cos(dw) {cs(dw, ref['dw']):.6f}{extra}{msg}", flush=True)
    except Exception:
        print(f"{name}: EXC", traceback.format_exc(), flush=True)

for p in sorted(glob.glob(os.path.join(ROOT, "tests/golden/*.npz"))):
    n = os.path.basename(p)[:-4]
    if n == "kat": continue
    g = dict(np.load(p))
    run(n, g["x"], g["w"], g["label"], float(g["s"]), float(g["m"]), bool(g["easy_margin"]), g)
for (B, D, C, chunk) in [(256, 512, 5000, 1024), (1024, 512, 20000, None), (512, 2048, 1000, None)]:
    gen = torch.Generator().manual_seed(B + C)
    w = ((torch.rand(C, D, generator=gen) * 2 - 1) * (6.0 / (C + D)) ** 0.5).numpy()
    x = torch.randn(B, D, generator=gen).numpy(); y = torch.randint(0, C, (B,), generator=gen).numpy()
    for easy in (False, True):
        ref = orc.head_forward_backward(x, w, y, 30.0, 0.5, easy)
        run(f"rand B{B} D{D} C{C} easy={easy}", x, w, y, 30.0, 0.5, easy, ref, chunk)
