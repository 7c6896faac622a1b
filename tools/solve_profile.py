#!/usr/bin/env python
"""One MG-PCG solve of the 3-D cantilever (reference production settings) for `ncu --metrics gpu__time_duration.sum`.
Usage: python tools/solve_profile.py [EX EY EZ] [levels]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import ndr_b200.pyVoxelFEM as V
FIX = os.path.join(ROOT, "tests", "golden", "fixtures")
grid = [int(a) for a in sys.argv[1:4]] if len(sys.argv) >= 4 else [128, 64, 64]
levels = int(sys.argv[4]) if len(sys.argv) >= 5 else 3
g = V.TensorProductSimulator([1, 1, 1], [np.zeros(3), np.array([2.0, 1.0, 1.0])], grid)
g.readMaterial(os.path.join(FIX, "materials", "B9Creator.material"))
g.setElementDensities(np.random.default_rng(88).uniform(0, 1, g.numElements()))
g.applyDisplacementsAndLoadsFromFile(os.path.join(FIX, "bcs", "3d", "cantilever_flexion.bc"))
g.E_0, g.E_min, g.gamma = 1.0, 1e-4, 3.0
mg = g.multigridSolver(levels)
f = torch.from_numpy(g.buildLoadVector()).cuda()
u0 = torch.zeros_like(f); out = torch.empty_like(f)
kw = dict(mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True, out=out)
reps = int(os.environ.get("REPS", "2"))
for r in range(reps):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    mg.preconditionedConjugateGradient(u0, f, 100, 1e-4, **kw)
    torch.cuda.synchronize()
    print("solve %d: %.2f ms, %d iterations, residual %.3e" % (r, (time.perf_counter() - t0) * 1e3, mg.lastIterations, mg.lastResidualNorm), flush=True)
