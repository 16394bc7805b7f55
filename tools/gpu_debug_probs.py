"""Debug aid for the stored-probabilities backward: compares dW / dX with the recompute path, tile by tile."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib

B, D, C = [int(v) for v in (sys.argv[1:4] if len(sys.argv) >= 4 else (1000, 512, 20000))]
g = torch.Generator().manual_seed(1)
w = ((torch.rand(C, D, generator=g) * 2 - 1) * 0.05)
x = torch.randn(B, D, generator=g)
y = torch.randint(0, C, (B,), generator=g)
res = {}
for name, flags in (("probs", _lib.KEEP_PROBS), ("recompute", 0)):
    head = ArcMarginProduct(D, C, kernel_flags=flags).cuda()
    with torch.no_grad():
        head.weight.copy_(w)
    xg = x.cuda().requires_grad_(True)
    loss = head.forward_loss(xg, y.cuda())
    loss.backward()
    torch.cuda.synchronize()
    res[name] = (loss.item(), xg.grad.clone(), head.weight.grad.clone())
p, r = res["probs"], res["recompute"]
print("loss", p[0], r[0])
dw_p, dw_r = p[2], r[2]
nan_rows = torch.isnan(dw_p).any(dim=1).nonzero().flatten()
print("NaN rows in dW(probs):", nan_rows.numel(), nan_rows[:10].tolist(), nan_rows[-10:].tolist())
print("NaN in dX(probs):", torch.isnan(p[1]).sum().item())
ok = ~torch.isnan(dw_p).any(dim=1)
cos = torch.nn.functional.cosine_similarity(dw_p[ok].flatten().double(), dw_r[ok].flatten().double(), dim=0)
print("cos dW (non-NaN rows)", cos.item(), "cos dX", torch.nn.functional.cosine_similarity(p[1].flatten().double(), r[1].flatten().double(), dim=0).item())
# per 128-class tile relative error
err = (dw_p - dw_r).norm(dim=1) / (dw_r.norm(dim=1) + 1e-30)
T = (C + 127) // 128
worst = []
for t in range(T):
    e = err[t * 128:(t + 1) * 128]
    e = e[~torch.isnan(e)]
    if e.numel():
        worst.append((e.max().item(), t))
worst.sort(reverse=True)
print("worst tiles (max row rel err, tile):", worst[:8])
print("median row rel err", err[~torch.isnan(err)].median().item())
# column halves
for h in range(2):
    sl = slice(h * 256, (h + 1) * 256)
    print("half", h, "cos", torch.nn.functional.cosine_similarity(dw_p[ok][:, sl].flatten().double(), dw_r[ok][:, sl].flatten().double(), dim=0).item())
# label rows
lab_rows = torch.unique(y)
print("label rows rel err median", err[lab_rows.cuda()].median().item(), "non-label median", err.median().item())
# reference WITHOUT the normalise-backward projection (compare with a run under ARCB_DEBUG_EPI=4: t_j forced to 0)
if os.environ.get("ARCB_DEBUG_EPI") == "4":
    import math
    xd, wd, yd = x.cuda().double(), w.cuda().double(), y.cuda()
    xh = torch.nn.functional.normalize(xd, dim=1); wn = wd.norm(dim=1).clamp_min(1e-12); wh = wd / wn[:, None]
    cos = xh @ wh.t()
    sine = torch.sqrt((1 - cos ** 2).clamp(0, 1)); phi = cos * math.cos(0.5) - sine * math.sin(0.5)
    th, mm = math.cos(math.pi - 0.5), math.sin(math.pi - 0.5) * 0.5
    hot = torch.zeros_like(cos).scatter_(1, yd.view(-1, 1), 1.0)
    sub = torch.where(cos > th, phi, cos - mm)
    z = 30.0 * (hot * sub + (1 - hot) * cos)
    p_ = torch.softmax(z, dim=1)
    dphi = torch.where(cos > th, math.cos(0.5) + math.sin(0.5) * cos / sine.clamp_min(1e-30), torch.ones_like(cos))
    g = (p_ - hot) * 30.0 / B * (hot * dphi + (1 - hot))
    dwh = g.t() @ xh
    ref_noproj = (dwh / wn[:, None]).float()
    okr = ~torch.isnan(dw_p).any(dim=1)
    print("no-projection check: cos", torch.nn.functional.cosine_similarity(dw_p[okr].flatten().double(), ref_noproj[okr].flatten().double(), dim=0).item(),
          "norm ratio", (dw_p[okr].norm() / ref_noproj[okr].norm()).item())
