"""Per-tile timeline of the persistent backward (ARCB_TRACE=1): where do the MMA issuer, the producer and
the epilogue spend their cycles?  Prints per-kind averages for CTAs 0..7."""
import os, sys
os.environ["ARCB_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np, torch
from ddp_classification_pytorch_b200 import _lib
from ddp_classification_pytorch_b200.head import CudaEngine
B, D, C = int(os.environ.get("TRACE_B", 1024)), int(os.environ.get("TRACE_D", 512)), int(os.environ.get("TRACE_C", 200000))
chunk = int(os.environ.get("ARCB_CHUNK_CLASSES", 256 * 29))
g = torch.Generator(device="cuda").manual_seed(0)
x = torch.randn(B, D, generator=g, device="cuda"); w = (torch.rand(C, D, generator=g, device="cuda") - 0.5) * 0.01
y = torch.randint(0, C, (B,), generator=g, device="cuda")
eng = CudaEngine(); L = _lib.lib()
cfg = {"s": 30.0, "m": 0.5, "easy_margin": False, "class_offset": 0, "chunk_classes": chunk}
stats, state = eng.shard_forward(x, w, y, cfg)
lse, zt, loss = eng.merge_loss(stats, 1, B)
Bq, Cs, Dq, ws, ws_bytes = state[:5]
Dp = L.arcb_padded_dim(D)
nbytes = L.arcb_scratch_bytes(B, 4 * chunk) + (1 << 20)
scratch = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
dw = torch.empty(C, D, device="cuda"); dxhat = torch.empty(B, Dp, device="cuda"); one = torch.ones((), device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(3):
    _lib.check(L.arcb_shard_backward(B, C, D, 30.0, ws.data_ptr(), ws_bytes, lse.data_ptr(), zt.data_ptr(), one.data_ptr(), None, 0,
                                     scratch.data_ptr(), nbytes, dw.data_ptr(), dxhat.data_ptr(), 0, st))
torch.cuda.synchronize()
tr = scratch[-(1 << 20):].view(torch.int64).cpu().numpy()[: 8 * 1024 * 12].reshape(8, 1024, 12)
names = {0: "G", 1: "X", 2: "W"}
for cta in (0, 2):
    t = tr[cta]; n = int((t[:, 2] > 0).sum()); t = t[:n]
    print(f"CTA {cta}: {n} tiles, span {(t[-1,6]-t[0,0])/1e3:.0f} kcyc")
    for k in (0, 1, 2):
        m = t[:, 9] == k
        if not m.any(): continue
        q = t[m]
        mma_total = (q[:, 2] - q[:, 0]).mean(); wait_tempty = (q[:, 1] - q[:, 0]).mean(); wait_full = q[:, 3].mean()
        epi_wait = (q[:, 5] - q[:, 4]).mean(); epi_work = (q[:, 6] - q[:, 5]).mean()
        prod_total = (q[:, 11] - q[:, 10]).mean(); prod_empty = q[:, 7].mean(); prod_flag = q[:, 8].mean()
        print(f"  {names[k]} x{int(m.sum()):4d}: MMA item {mma_total:7.0f} (wait tempty {wait_tempty:6.0f}, wait full {wait_full:6.0f}) | "
              f"epilogue wait tfull {epi_wait:7.0f} work {epi_work:6.0f} | producer item {prod_total:7.0f} (wait empty {prod_empty:6.0f}, flag {prod_flag:6.0f})")
    # steady-state sample of consecutive tiles
    print("   first 12 tiles: kind, mma_start, mma_end-start, epi_work:", [(names[int(r[9])], int(r[0]-t[0,0]), int(r[2]-r[0]), int(r[6]-r[5])) for r in t[20:32]])

print("per-CTA totals (kcycles): producer flag wait | producer empty wait | MMA full wait | MMA tempty wait | epilogue tfull wait | epilogue work | span")
for cta in range(8):
    t = tr[cta]; n = int((t[:, 6] > 0).sum()); t = t[:n]
    if n == 0: continue
    span = (t[-1, 6] - t[0, 4]) / 1e3
    print(f"  CTA {cta}: n={n:4d} flag {t[:,8].sum()/1e3:7.0f} | empty {t[:,7].sum()/1e3:7.0f} | full {t[:,3].sum()/1e3:7.0f} | "
          f"tempty {(t[:,1]-t[:,0]).clip(min=0).sum()/1e3:7.0f} | tfull {(t[:,5]-t[:,4]).sum()/1e3:7.0f} | work {(t[:,6]-t[:,5]).sum()/1e3:7.0f} | span {span:7.0f}")
