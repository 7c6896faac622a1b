#!/usr/bin/env python
"""torchrun check of the NCCL slab solver against the single-GPU solver on the same global problem.
   python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/slab_check.py [EX EY EZ levels]"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from ndr_b200 import _lib, slab as S
import ndr_b200.pyVoxelFEM as V
from bench import layer_densities, MATERIAL, BC
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); _lib.check(_lib.lib().vfem_set_device(local))
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
grid = [int(a) for a in sys.argv[1:4]] if len(sys.argv) >= 5 else [64 * world, 32, 32]
levels = int(sys.argv[4]) if len(sys.argv) >= 5 else 3
bbmin, bbmax = np.zeros(3), np.array([2.0, 1.0, 1.0])
pe = grid[1] * grid[2]
kw = dict(mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True, return_info=True)
comm = S.make_nccl_comm(dist, rank, world)
sl = S.SlabSolver(grid, bbmin, bbmax, levels, rank, world, comm, S.dist_host_sum(dist, torch.device("cuda", local)), MATERIAL)
lay = sl.layout
sl.tps.setElementDensities(layer_densities(lay.layer_lo, lay.layer_hi, pe))
sl.apply_bc_file(BC)
sl.tps.E_0, sl.tps.E_min, sl.tps.gamma = 1.0, 1e-4, 3.0
mg = sl.multigrid()
fl = sl.tps.buildLoadVector()
u, its, res = mg.preconditionedConjugateGradient(np.zeros_like(fl), fl, 100, 1e-10, **kw)
# single-GPU solve of the global problem on every rank (small grid) as the comparison
g = V.TensorProductSimulator([1, 1, 1], [bbmin, bbmax], grid)
g.readMaterial(MATERIAL); g.setElementDensities(layer_densities(0, grid[0], pe)); g.applyDisplacementsAndLoadsFromFile(BC)
g.E_0, g.E_min, g.gamma = 1.0, 1e-4, 3.0
f = g.buildLoadVector()
ug, itg, _ = g.multigridSolver(levels).preconditionedConjugateGradient(np.zeros_like(f), f, 100, 1e-10, **kw)
err = np.max(np.abs(u[lay.owned_node_rows()] - ug[lay.owned_global_node_rows()])) / np.max(np.abs(ug))
print("rank %d/%d: slab iterations %d (single %d), residual %.3e, owned-plane rel err vs single-GPU solve %.3e" % (rank, world, its, itg, res, err), flush=True)
assert err < 1e-9 and its == itg
dist.barrier(); dist.destroy_process_group()
