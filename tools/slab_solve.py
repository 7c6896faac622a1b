#!/usr/bin/env python
"""Times the slab-decomposed MG-PCG solve (reference production settings) and the halo-exchanged K-apply of a 3-D
cantilever on N GPUs.  python -m torch.distributed.run --nproc-per-node N ... tools/slab_solve.py EX EY EZ levels [reps]"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from ndr_b200 import _lib, slab as S
from ndr_b200.voxelfem import _dev_ptr
from bench import layer_densities, MATERIAL, BC
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local); _lib.check(_lib.lib().vfem_set_device(local))
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
grid = [int(a) for a in sys.argv[1:4]]; levels = int(sys.argv[4]); reps = int(sys.argv[5]) if len(sys.argv) > 5 else 3
bbmin, bbmax = np.zeros(3), np.array([2.0, 1.0, 1.0])
comm = S.make_nccl_comm(dist, rank, world)
sl = S.SlabSolver(grid, bbmin, bbmax, levels, rank, world, comm, S.dist_host_sum(dist, torch.device("cuda", local)), MATERIAL)
lay = sl.layout
sl.tps.setElementDensities(layer_densities(lay.layer_lo, lay.layer_hi, lay.plane_elems))
sl.apply_bc_file(BC)
sl.tps.E_0, sl.tps.E_min, sl.tps.gamma = 1.0, 1e-4, 3.0
mg = sl.multigrid()
f = torch.from_numpy(sl.tps.buildLoadVector()).cuda()
u0 = torch.zeros_like(f); out = torch.empty_like(f)
kw = dict(mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True, out=out)
nn = (grid[0] + 1) * (grid[1] + 1) * (grid[2] + 1)
def tmax(ms):
    t = torch.tensor([ms], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); return float(t.item())
for r in range(reps):
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); mg.preconditionedConjugateGradient(u0, f, 100, 1e-4, **kw); e1.record(); torch.cuda.synchronize()
    ms = tmax(e0.elapsed_time(e1))
    if rank == 0:
        print("solve %d: %.2f ms on %d GPUs, %d iterations, residual %.3e, %d dofs, %.1f Mdof/s" % (r, ms, world, mg.lastIterations, mg.lastResidualNorm, 3 * nn, 3 * nn / ms / 1e3), flush=True)
J = 0.5 * sl.dot(f, out)
u = torch.randn_like(f); Ku = torch.empty_like(f)
L = _lib.lib()
def step():
    _lib.check(L.vfem_mg_apply_k_halo(mg._h, _dev_ptr(u, u.numel()), _dev_ptr(Ku, Ku.numel()), 1))
for _ in range(5): step()
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): step()
e1.record(); torch.cuda.synchronize()
ms = tmax(e0.elapsed_time(e1) / 20)
if rank == 0:
    print("compliance %.9g; K-apply + halo exchange: %.3f ms -> %.1f GDOF/s over %d GPUs (%.1f%% of %d x HBM roofline)" % (
        J, ms, 3 * nn / ms / 1e6, world, 100 * 8 * (6 * nn + grid[0] * grid[1] * grid[2]) / (ms * 1e-3) / (world * 6530.6e9), world), flush=True)
if os.environ.get("OPS"):  # per-operator device times (collective: every rank makes the same calls)
    import ctypes as C
    names = {0: "GS sweep", 1: "residual", 2: "restrict", 3: "prolong", 7: "masked apply"}
    msv = C.c_double()
    def t(op, l, n=20):
        _lib.check(L.vfem_mg_time_op(mg._h, op, l, n, C.byref(msv))); return msv.value * 1e3
    for l in range(levels):
        line = "level %d: " % l + "  ".join("%s %.1f us" % (names[o], t(o, l)) for o in (0, 1, 2, 3, 7))
        if rank == 0: print(line, flush=True)
    line = "coarse solve %.1f us   refresh %.1f us   FMG preconditioner pass %.1f us" % (t(4, levels), t(5, 0, 5), t(6, 0, 10))
    if rank == 0: print(line, flush=True)
dist.barrier(); dist.destroy_process_group()
