import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_samples_b200 as prim
N = 1 << 28
src = torch.empty(N, dtype=torch.int32, device="cuda"); dst = torch.empty(N, dtype=torch.int32, device="cuda")
prim.generate(src, "u31", 0, N, 2009)
def t(fn, it=20):
    for _ in range(4): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(it): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / it * 1e3
print("scan 1 x 2^28        %.1f us" % t(lambda: prim.scanExclusiveLarge(dst, src, 1, N)))
print("scan 1024 x 2^18     %.1f us" % t(lambda: prim.scanExclusiveLarge(dst, src, 1024, 1 << 18)))
print("scan 16384 x 16384   %.1f us" % t(lambda: prim.scanExclusiveLarge(dst, src, 16384, 16384)))
cnt = torch.zeros(1, dtype=torch.int32, device="cuda")
prim.generate(src, "key_mod", 0, N, 7, 5)
print("compact 80%%          %.1f us" % t(lambda: prim.filter_arr(dst, cnt, src, N)))
src.neg_()
print("compact 0%%           %.1f us" % t(lambda: prim.filter_arr(dst, cnt, src, N)))
