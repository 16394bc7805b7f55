"""First-contact diagnostics for the tcgen05 GEMM core; prints per-config errors (not a test)."""
import sys, os, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ddp_classification_pytorch_b200 import _lib
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from test_gpu_gemm import _run

print(torch.cuda.get_device_name(0), torch.version.cuda, flush=True)
for (M, N, K) in [(128, 256, 64), (128, 256, 512), (256, 512, 192), (100, 300, 72), (1024, 1024, 512)]:
    for a_mn in (False, True):
        for b_mn in (False, True):
            try:
                e = _run(M, N, K, a_mn, b_mn, 1)
                print(f"M{M} N{N} K{K} a_mn={int(a_mn)} b_mn={int(b_mn)} relerr={e:.3e}", flush=True)
            except Exception as ex:
                print(f"M{M} N{N} K{K} a_mn={int(a_mn)} b_mn={int(b_mn)} EXC {ex}", flush=True)
for split in (3, 8):
    for mn in (False, True):
        try:
            print("split", split, mn, _run(256, 512, 64 * 37, mn, mn, split), flush=True)
        except Exception as ex:
            print("split EXC", ex, flush=True)
