#!/usr/bin/env python
"""MG-PCG solves of the 3-D cantilever (reference production settings) + per-operator device timings (vfem_mg_time_op).
Usage: python tools/solve_profile.py [EX EY EZ] [levels]"""
import ctypes as C, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import ndr_b200.pyVoxelFEM as V
from ndr_b200._lib import lib, check
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
if not os.environ.get("NO_OPS"):
    names = {0: "GS sweep", 1: "residual", 2: "restrict", 3: "prolong", 7: "masked apply"}
    ms = C.c_double()
    def t(op, l, n=20):
        check(lib().vfem_mg_time_op(mg._h, op, l, n, C.byref(ms))); return ms.value * 1e3
    for l in range(levels):
        print("level %d: " % l + "  ".join("%s %.1f us" % (names[o], t(o, l)) for o in (0, 1, 2, 3, 7)), flush=True)
    print("coarse solve %.1f us   refresh %.1f us   FMG preconditioner pass %.1f us" % (t(4, levels), t(5, 0, 5), t(6, 0, 10)), flush=True)
