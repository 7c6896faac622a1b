"""host-array K-apply (the e2e leg of bench.py): sweep of the H2D / kernel / D2H pipeline depth + the raw copy rates"""
import os, sys, time, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
if len(sys.argv) > 1 and sys.argv[1] == "child":
    import numpy as np, torch
    import ndr_b200.pyVoxelFEM as V
    from bench import MATERIAL, BC
    g = V.TensorProductSimulator([1, 1, 1], [np.zeros(3), np.array([2.0, 1.0, 1.0])], [128, 64, 64])
    g.readMaterial(MATERIAL); g.setElementDensities(np.random.default_rng(88).uniform(0, 1, g.numElements())); g.applyDisplacementsAndLoadsFromFile(BC)
    nn = g.numNodes()
    uh = torch.randn(nn, 3, dtype=torch.float64).pin_memory().numpy()
    oh = torch.empty(nn, 3, dtype=torch.float64).pin_memory().numpy()
    for _ in range(5): g.applyK(uh, out=oh)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(50): g.applyK(uh, out=oh)
    torch.cuda.synchronize(); ms = (time.perf_counter() - t0) / 50 * 1e3
    print("VFEM_HOST_CHUNKS=%s: %.3f ms per host apply (%.2f GDOF/s)" % (os.environ.get("VFEM_HOST_CHUNKS", "default"), ms, 3 * nn / ms / 1e6))
    if os.environ.get("VFEM_HOST_CHUNKS") == "1":
        a = torch.from_numpy(uh); d = torch.empty(nn, 3, dtype=torch.float64, device="cuda"); b = torch.from_numpy(oh)
        for name, fn in (("H2D", lambda: d.copy_(a, non_blocking=True)), ("D2H", lambda: b.copy_(d, non_blocking=True))):
            fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
            for _ in range(20): fn()
            torch.cuda.synchronize(); ms = (time.perf_counter() - t0) / 20 * 1e3
            print("  raw %s of %.1f MB pinned: %.3f ms = %.1f GB/s" % (name, nn * 24 / 1e6, ms, nn * 24 / ms / 1e6))
else:
    for c in ("1", "2", "4", "6", "8", "12", "16"):
        subprocess.run([sys.executable, __file__, "child"], env=dict(os.environ, VFEM_HOST_CHUNKS=c))
