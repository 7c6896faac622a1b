"""Slab decomposition of the fine-level K-apply across GPUs.

The grid is cut into contiguous slabs along storage axis 0 (the slowest index, so every node plane is one
contiguous range).  Rank r owns element layers [e0, e1) and the node planes e0..e1; the interface plane is
replicated on both neighbours.  Each rank's local simulator additionally carries ONE ghost element layer per
interior face, so the gather-form apply computes complete rows for every owned node once the ghost node
plane holds the neighbour's values.  Per apply that is one node plane (n1*n2*N doubles) sent and received
per interior face - `torch.distributed` send/recv over NCCL (NVLink) on GPU, gloo in the CPU tests.
"""
import numpy as np


def partition(E0, world, rank):
    """Element-layer range [e0, e1) owned by `rank`; layers are split evenly (remainder to the low ranks)."""
    base, rem = divmod(E0, world)
    e0 = rank * base + min(rank, rem)
    return e0, e0 + base + (1 if rank < rem else 0)


class SlabLayout:
    def __init__(self, global_elems, bbmin, bbmax, rank, world):
        self.rank, self.world = rank, world
        self.gE = [int(e) for e in global_elems]
        self.N = len(self.gE)
        self.e0, self.e1 = partition(self.gE[0], world, rank)
        self.ghost_lo = 1 if rank > 0 else 0
        self.ghost_hi = 1 if rank < world - 1 else 0
        lo, hi = self.e0 - self.ghost_lo, self.e1 + self.ghost_hi
        self.local_elems = [hi - lo] + self.gE[1:]
        bbmin, bbmax = np.asarray(bbmin, float), np.asarray(bbmax, float)
        h0 = (bbmax[0] - bbmin[0]) / self.gE[0]
        self.local_bbmin = bbmin.copy(); self.local_bbmax = bbmax.copy()
        self.local_bbmin[0] = bbmin[0] + lo * h0
        self.local_bbmax[0] = bbmin[0] + hi * h0
        self.elem_lo, self.elem_hi = lo, hi
        self.plane_nodes = int(np.prod([e + 1 for e in self.gE[1:]]))
        self.local_planes = hi - lo + 1
        # local plane indices
        self.first_owned = self.ghost_lo                    # global plane e0
        self.last_owned = self.local_planes - 1 - self.ghost_hi  # global plane e1

    def local_node_slice(self):
        """Global node-plane range [p0, p1] covered by the local (ghost-extended) grid."""
        return self.elem_lo, self.elem_hi

    def owned_weight_planes(self):
        """Planes this rank counts in global reductions: the interface plane belongs to the lower rank."""
        first = self.first_owned + (1 if self.rank > 0 else 0)
        return first, self.last_owned


def exchange_ghost_planes(u_planes, layout, dist):
    """u_planes: torch tensor viewed as (local_planes, plane_nodes*N).  Fills the ghost planes with the
    neighbours' first interior planes (one send + one recv per interior face)."""
    ops = []
    L = layout
    if L.ghost_lo:
        ops.append(dist.P2POp(dist.isend, u_planes[L.first_owned + 1].contiguous(), L.rank - 1))
        ops.append(dist.P2POp(dist.irecv, u_planes[0], L.rank - 1))
    if L.ghost_hi:
        ops.append(dist.P2POp(dist.isend, u_planes[L.last_owned - 1].contiguous(), L.rank + 1))
        ops.append(dist.P2POp(dist.irecv, u_planes[L.local_planes - 1], L.rank + 1))
    if ops:
        for r in dist.batch_isend_irecv(ops):
            r.wait()


class SlabKApply:
    """Distributed matrix-free K(rho) u: ghost exchange + the local CUDA apply through the C ABI."""

    def __init__(self, global_elems, bbmin, bbmax, rank, world, material_path, rho_local):
        from . import pyVoxelFEM as V
        self.layout = L = SlabLayout(global_elems, bbmin, bbmax, rank, world)
        self.tps = V.TensorProductSimulator([1] * L.N, [L.local_bbmin, L.local_bbmax], L.local_elems)
        self.tps.readMaterial(material_path)
        self.tps.setElementDensities(rho_local)

    def apply(self, u, out, dist=None):
        L = self.layout
        if L.world > 1:
            exchange_ghost_planes(u.view(L.local_planes, -1), L, dist)
        self.tps.applyK(u, out=out)
        return out
