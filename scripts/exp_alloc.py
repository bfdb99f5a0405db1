"""How long do cudaMalloc / cudaFree / first touch of large blocks take on this box, cold and warm?  (run under gpurun)"""
import ctypes as C
import time
import torch

rt = C.CDLL("libcudart.so.12") if False else None
torch.cuda.init()
t0 = time.perf_counter()
torch.zeros(1, device="cuda")
torch.cuda.synchronize()
print("context %.3f s" % (time.perf_counter() - t0))
import glob, os
cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
rt = C.CDLL(cands[0])
rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
rt.cudaFree.argtypes = [C.c_void_p]
rt.cudaMemset.argtypes = [C.c_void_p, C.c_int, C.c_size_t]
GB = 1 << 30
for rnd in range(3):
    ptrs = []
    t0 = time.perf_counter()
    for i in range(16):
        p = C.c_void_p()
        assert rt.cudaMalloc(C.byref(p), 2 * GB) == 0
        ptrs.append(p)
    t1 = time.perf_counter()
    for p in ptrs:
        rt.cudaMemset(p, 0, 2 * GB)
    rt.cudaDeviceSynchronize()
    t2 = time.perf_counter()
    for p in ptrs:
        rt.cudaFree(p)
    t3 = time.perf_counter()
    print("round %d: malloc 32 GB %.3f s, memset %.3f s, free %.3f s" % (rnd, t1 - t0, t2 - t1, t3 - t2))
# pinned host memory
for rnd in range(2):
    t0 = time.perf_counter()
    h = torch.empty(2 * GB // 4, dtype=torch.float32).pin_memory()
    t1 = time.perf_counter()
    d = torch.empty_like(h, device="cuda")
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    t3 = time.perf_counter()
    print("pinned 2 GB: alloc+pin %.3f s, H2D %.3f s (%.1f GB/s)" % (t1 - t0, t3 - t2, 2.147 / (t3 - t2)))
    del h, d
