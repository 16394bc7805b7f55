"""Debug aid: dW kernel of the stored-probabilities backward under ARCB_DEBUG_EPI=5/6/7 (constant / raw accumulator / t_j)."""
import math, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
from ddp_classification_pytorch_b200 import ArcMarginProduct, _lib
mode = int(os.environ.get("ARCB_DEBUG_EPI", "0"))
B, D, C = 1000, 512, 20000
g = torch.Generator().manual_seed(1)
w = ((torch.rand(C, D, generator=g) * 2 - 1) * 0.05); x = torch.randn(B, D, generator=g); y = torch.randint(0, C, (B,), generator=g)
head = ArcMarginProduct(D, C, kernel_flags=_lib.KEEP_PROBS).cuda()
with torch.no_grad(): head.weight.copy_(w)
xg = x.cuda().requires_grad_(True)
loss = head.forward_loss(xg, y.cuda()); loss.backward(); torch.cuda.synchronize()
dw = head.weight.grad
print(f"mode {mode}: loss {loss.item():.5f}  NaN rows {torch.isnan(dw).any(1).sum().item()}  zero rows {(dw == 0).all(1).sum().item()}  "
      f"rows==1 {(dw == 1).all(1).sum().item()}  mean|dw| {dw[~torch.isnan(dw).any(1)].abs().mean().item():.3e}")
xd, wd, yd = x.cuda().double(), w.cuda().double(), y.cuda()
xh = torch.nn.functional.normalize(xd, dim=1); wn = wd.norm(dim=1); wh = wd / wn[:, None]
cos = xh @ wh.t(); sine = torch.sqrt((1 - cos ** 2).clamp(0, 1)); phi = cos * math.cos(0.5) - sine * math.sin(0.5)
th, mm = math.cos(math.pi - 0.5), math.sin(math.pi - 0.5) * 0.5
hot = torch.zeros_like(cos).scatter_(1, yd.view(-1, 1), 1.0)
z = 30.0 * (hot * torch.where(cos > th, phi, cos - mm) + (1 - hot) * cos)
p_ = torch.softmax(z, dim=1)
dphi = torch.where(cos > th, math.cos(0.5) + math.sin(0.5) * cos / sine.clamp_min(1e-30), torch.ones_like(cos))
G = (p_ - hot) * 30.0 / B * (hot * dphi + (1 - hot))
dwh = G.t() @ xh                       # raw accumulator the kernel should hold
tj = (wh * dwh).sum(1)                 # projection
def cmp(name, got, ref):
    ok = ~torch.isnan(got).any(1)
    c = torch.nn.functional.cosine_similarity(got[ok].double().flatten(), ref[ok].flatten(), dim=0).item()
    print(f"   vs {name}: cos {c:.6f}  norm ratio {(got[ok].double().norm() / ref[ok].norm()).item():.4f}")
    for t in (0, 1, 73, 74, 75, 147, 148, 155, 156):
        sl = slice(t * 128, min(C, (t + 1) * 128)); gg, rr = got[sl].double(), ref[sl]
        print(f"      tile {t}: nan {torch.isnan(gg).any().item()}  cos {torch.nn.functional.cosine_similarity(torch.nan_to_num(gg).flatten(), rr.flatten(), dim=0).item():.4f}  |got| {torch.nan_to_num(gg).norm().item():.3e} |ref| {rr.norm().item():.3e}")
def chunks(name, got, ref, t):
    sl = slice(t * 128, min(C, (t + 1) * 128)); gg, rr = torch.nan_to_num(got[sl].double()), ref[sl]
    cs = [torch.nn.functional.cosine_similarity(gg[:, c:c + 32].flatten(), rr[:, c:c + 32].flatten(), dim=0).item() for c in range(0, D, 32)]
    rs = [torch.nn.functional.cosine_similarity(gg[r:r + 16].flatten(), rr[r:r + 16].flatten(), dim=0).item() for r in range(0, gg.shape[0], 16)]
    print(f"   {name} tile {t}: cos per 32-col chunk", [round(c, 2) for c in cs])
    print(f"   {name} tile {t}: cos per 16-row group", [round(c, 2) for c in rs])
    # rows with a label: is the row a multiple of SOME x_hat row?
    rows_l = [int(j) for j in torch.unique(yd) if t * 128 <= int(j) < (t + 1) * 128][:6]
    for j in rows_l:
        gvec = torch.nan_to_num(got[j].double()); sims = (xh @ gvec) / (gvec.norm() + 1e-30)
        best = int(sims.abs().argmax()); want = (yd == j).nonzero().flatten().tolist()
        print(f"      class {j}: labelled samples {want}, row is most parallel to x_hat[{best}] (|cos| {sims[best].abs().item():.3f}); cos with reference row {torch.nn.functional.cosine_similarity(gvec, ref[j], dim=0).item():.3f}")
if mode in (6, 8):
    cmp("raw accumulator G^T x_hat", dw, dwh)
    for t in (0, 1, 74): chunks("raw acc", dw, dwh, t)
    # which reference 64-column block (of which tile) does each 64-column block of tile 0 / 74 hold?
    for t in (0, 74):
        gg = torch.nan_to_num(dw[t * 128:(t + 1) * 128].double(), posinf=0.0, neginf=0.0)
        for b in range(8):
            gb = gg[:, b * 64:(b + 1) * 64]
            best = (0.0, None)
            for t2 in (t, t + 1, (t + 74) % 156, max(t - 74, 0)):
                rr = dwh[t2 * 128:(t2 + 1) * 128]
                for b2 in range(8):
                    c = torch.nn.functional.cosine_similarity(gb.flatten(), rr[:, b2 * 64:(b2 + 1) * 64].flatten(), dim=0).item()
                    if abs(c) > abs(best[0]): best = (c, (t2, b2))
            print(f"   tile {t} block {b}: best match ref (tile, block) {best[1]} cos {best[0]:.3f}  |got| {gb.norm().item():.3e}")
if mode == 7: cmp("t_j broadcast", dw, tj[:, None].expand(-1, D))
if mode in (0, 4):
    ref = ((dwh - (tj[:, None] * wh if mode == 0 else 0)) / wn[:, None])
    cmp("final dW" if mode == 0 else "dW without projection", dw, ref)
    for t in (0, 74): chunks("dW", dw, ref, t)
