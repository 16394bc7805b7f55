"""Probe: does cuTensorMapEncodeTiled accept a rank-3 map whose second stride is smaller than the first?"""
import ctypes, torch
cuda = ctypes.CDLL("libcuda.so.1")
t = torch.zeros(1024, 512, dtype=torch.bfloat16, device="cuda")
tm = (ctypes.c_uint8 * 128)()
def enc(dims, strides, box):
    r = len(dims)
    D = (ctypes.c_uint64 * r)(*dims); S = (ctypes.c_uint64 * (r - 1))(*strides)
    Bx = (ctypes.c_uint32 * r)(*box); E = (ctypes.c_uint32 * r)(*([1] * r))
    return cuda.cuTensorMapEncodeTiled(ctypes.byref(tm), 9, r, ctypes.c_void_p(t.data_ptr()), D, S, Bx, E, 0, 3, 2, 0)
print("2D            ", enc([512, 1024], [1024], [64, 64]))
print("3D {64,K,blk} strides {pitch,128}", enc([64, 1024, 8], [1024, 128], [64, 64, 4]))
print("3D {64,blk,K} strides {128,pitch}", enc([64, 8, 1024], [128, 1024], [64, 4, 64]))
