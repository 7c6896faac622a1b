#!/usr/bin/env python
"""Host-array K-apply (vfem_tps_apply_k with VFEM_HOST pointers): time per call for the chunk count in VFEM_HOST_CHUNKS.
Usage: VFEM_HOST_CHUNKS=8 python tools/e2e_probe.py [EX EY EZ]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import ndr_b200.pyVoxelFEM as V
grid = [int(a) for a in sys.argv[1:4]] if len(sys.argv) >= 4 else [128, 64, 64]
g = V.TensorProductSimulator([1, 1, 1], [np.zeros(3), np.array([2.0, 1.0, 1.0])], grid)
g.setElementDensities(np.random.default_rng(88).uniform(0, 1, g.numElements()))
nn = g.numNodes()
uh = torch.randn(nn, 3, dtype=torch.float64).pin_memory().numpy()
oh = torch.empty(nn, 3, dtype=torch.float64).pin_memory().numpy()
for _ in range(5): g.applyK(uh, out=oh)
torch.cuda.synchronize(); t0 = time.perf_counter()
n = 50
for _ in range(n): g.applyK(uh, out=oh)
torch.cuda.synchronize(); ms = (time.perf_counter() - t0) / n * 1e3
ud = torch.from_numpy(uh).cuda(); od = torch.empty_like(ud); g.applyK(ud, out=od); ref = od.cpu().numpy()
print("chunks %s: %.3f ms/call, %.2f GDOF/s, %.1f GB/s each way, max |diff| vs device path %.1e" % (
    os.environ.get("VFEM_HOST_CHUNKS", "4"), ms, 3 * nn / ms / 1e6, uh.nbytes / ms / 1e6, np.abs(ref - oh).max()))
