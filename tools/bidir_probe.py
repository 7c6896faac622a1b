"""are H2D and D2H copies of pinned arrays concurrent on this box?  (the host-array K-apply pipelines them)"""
import time, torch
for mb in (13, 130):
    n = mb * 1000 * 1000 // 8
    a = torch.randn(n, dtype=torch.float64).pin_memory(); b = torch.empty(n, dtype=torch.float64).pin_memory()
    d0 = torch.empty(n, dtype=torch.float64, device="cuda"); d1 = torch.randn(n, dtype=torch.float64, device="cuda")
    s0, s1 = torch.cuda.Stream(), torch.cuda.Stream()
    def h2d():
        with torch.cuda.stream(s0): d0.copy_(a, non_blocking=True)
    def d2h():
        with torch.cuda.stream(s1): b.copy_(d1, non_blocking=True)
    for name, fns in (("H2D", (h2d,)), ("D2H", (d2h,)), ("both", (h2d, d2h))):
        for f in fns: f()
        torch.cuda.synchronize(); t0 = time.perf_counter()
        for _ in range(20):
            for f in fns: f()
        torch.cuda.synchronize(); ms = (time.perf_counter() - t0) / 20 * 1e3
        print("%4d MB %5s: %.3f ms  (%.1f GB/s per direction)" % (mb, name, ms, mb / ms))
