"""Slab decomposition of the elasticity solve across GPUs - host-side logic.

The grid is cut along storage axis 0 (the slowest index: every node plane is one contiguous range).  For a solver
with L coarsening levels the coarsest-level element layers are split evenly over the ranks; rank r owns the coarsest
layers [c0, c1) and, on level l, the node planes [c0, c1) * 2^(L-l) (the last rank also owns the final plane).  Its
local grids hold the coarsest layers [c0 - [r > 0], c1) refined 2^(L-l) times, so local fine = 2 x local coarse on every
level: 2^(L-l) ghost node planes below the owned range (only the nearest is ever read) and one ghost plane above it.

Per operator application one node plane per interior face crosses NVLink (ncclSend/ncclRecv enqueued by the C++ solver
on its own stream); dot products are one-scalar all-reduces; the coarsest level is gathered and solved redundantly on
every rank.  `torch.distributed` (NCCL on the GPUs, gloo in the CPU tests) is the control plane: it broadcasts the NCCL
unique id and reduces the node counts of force regions.  There is no reference counterpart (the reference is TBB on
one node); results are those of the undecomposed grid to roundoff and are asserted against the CPU restatement in the tests.
"""
import ctypes as C

import numpy as np


def partition(E0, world, rank):
    """Layer range [e0, e1) owned by `rank`; layers are split evenly (remainder to the low ranks)."""
    base, rem = divmod(E0, world)
    e0 = rank * base + min(rank, rem)
    return e0, e0 + base + (1 if rank < rem else 0)


class SlabLayout:
    """Index bookkeeping of one rank (pure Python / numpy: exercised on CPU with gloo)."""

    def __init__(self, global_elems, levels, rank, world):
        self.rank, self.world, self.levels = rank, world, levels
        self.gE = [int(e) for e in global_elems]
        self.N = len(self.gE)
        w = 1 << levels
        if self.gE[0] % w:
            raise RuntimeError("slab solve: axis-0 layers must be a multiple of 2^levels")
        EL = self.gE[0] // w
        if EL < world:
            raise RuntimeError("slab solve: %d coarsest layers cannot be split over %d ranks" % (EL, world))
        self.c0, self.c1 = partition(EL, world, rank)
        self.ghost_lo = w if rank > 0 else 0                    # fine ghost node planes below the owned range
        self.layer_lo = self.c0 * w - self.ghost_lo             # first local fine element layer (global index)
        self.layer_hi = self.c1 * w                             # one past the last local fine element layer
        self.local_elems = [self.layer_hi - self.layer_lo] + self.gE[1:]
        self.plane_nodes = int(np.prod([e + 1 for e in self.gE[1:]]))
        self.plane_elems = int(np.prod(self.gE[1:]))
        self.local_planes = self.local_elems[0] + 1
        self.own0 = self.ghost_lo                               # owned node planes (local, inclusive)
        self.own1 = self.local_planes - 1 if rank == world - 1 else self.local_planes - 2
        self.elem_own0 = self.ghost_lo                          # owned element layers (local, half-open)
        self.elem_own1 = self.local_elems[0]

    # ---- global <-> local array slicing (node fields are (planes*plane_nodes, N), element fields (layers*plane_elems,))
    def node_slice(self):
        """rows of a GLOBAL node field covered by the local grid (ghost planes included)"""
        return slice(self.layer_lo * self.plane_nodes, (self.layer_hi + 1) * self.plane_nodes)

    def elem_slice(self):
        return slice(self.layer_lo * self.plane_elems, self.layer_hi * self.plane_elems)

    def owned_node_rows(self):
        """rows of the LOCAL node field this rank owns"""
        return slice(self.own0 * self.plane_nodes, (self.own1 + 1) * self.plane_nodes)

    def owned_global_node_rows(self):
        g0 = self.layer_lo + self.own0
        return slice(g0 * self.plane_nodes, (g0 + self.own1 - self.own0 + 1) * self.plane_nodes)

    def owned_elem_rows(self):
        return slice(self.elem_own0 * self.plane_elems, self.elem_own1 * self.plane_elems)

    def owned_global_elem_rows(self):
        g0 = self.layer_lo + self.elem_own0
        return slice(g0 * self.plane_elems, (g0 + self.elem_own1 - self.elem_own0) * self.plane_elems)


def exchange_ghost_planes(v, layout, dist, group=None):
    """Reference (host-driven) halo exchange of a local node field through torch.distributed: the nearest ghost plane
    below / above the owned range <- the neighbour's last / first owned plane.  The C++ solver does the same with
    ncclSend/ncclRecv on its stream; this version is what the gloo tests and the K-apply-only path use."""
    L = layout
    planes = v.view(L.local_planes, -1)
    ops = []
    if L.rank > 0:
        ops.append(dist.P2POp(dist.isend, planes[L.own0].contiguous(), L.rank - 1, group))
        ops.append(dist.P2POp(dist.irecv, planes[L.own0 - 1], L.rank - 1, group))
    if L.rank < L.world - 1:
        ops.append(dist.P2POp(dist.isend, planes[L.own1].contiguous(), L.rank + 1, group))
        ops.append(dist.P2POp(dist.irecv, planes[L.own1 + 1], L.rank + 1, group))
    if ops:
        for r in dist.batch_isend_irecv(ops):
            r.wait()


class SlabSolver:
    """One rank of the slab-decomposed simulator + MG-PCG solver (CUDA, through the C ABI).

    comm: a `vfem_comm_t` handle (make_nccl_comm / make_thread_comms).  host_sum(int) -> int sums an integer over the
    ranks on the host (torch.distributed all_reduce, or a threading reduction in the single-GPU tests)."""

    def __init__(self, global_elems, bbmin, bbmax, levels, rank, world, comm, host_sum, material_path=None):
        from . import pyVoxelFEM as V
        from ._lib import check, lib
        from .voxelfem import _np, _ptr
        self.layout = L = SlabLayout(global_elems, levels, rank, world)
        self._comm, self._host_sum, self.levels = comm, host_sum, levels
        tps = V.TensorProductSimulator.__new__(V.TensorProductSimulator)
        deg = _np([1] * L.N, np.uint64)
        el = _np(L.gE, np.uint64)
        tps._bbmin, tps._bbmax, tps.N = _np(bbmin), _np(bbmax), L.N
        tps._h = C.c_void_p()
        check(lib().vfem_tps_create_slab(L.N, _ptr(deg), _ptr(tps._bbmin), _ptr(tps._bbmax), _ptr(el), L.layer_lo, L.layer_hi, C.byref(tps._h)))
        tps._elems = list(L.local_elems)
        self.tps = tps
        if material_path:
            tps.readMaterial(material_path)
        self.mg = None

    def apply_bc_file(self, path):
        """applyDisplacementsAndLoadsFromFile on the GLOBAL grid (TPS.hh:358-409): a force is split over the nodes of its
        region counted over all ranks (each node counted by its owner)."""
        from ._lib import check, lib
        from .voxelfem import _ptr, parse_bc_regions
        L, t = self.layout, self.tps
        for typ, lo, hi, comp, val in parse_bc_regions(path, t._bbmin, t._bbmax, t.N):
            if typ == "force":
                cnt = C.c_size_t(0)
                check(lib().vfem_tps_count_region_nodes(t._h, _ptr(lo), _ptr(hi), L.own0, L.own1, C.byref(cnt)))
                total = int(self._host_sum(int(cnt.value)))
                if total == 0:
                    raise RuntimeError("Force constraint region unmatched")
                check(lib().vfem_tps_add_force_region_split(t._h, _ptr(lo), _ptr(hi), _ptr(val), total))
            else:
                check(lib().vfem_tps_add_dirichlet_region(t._h, _ptr(lo), _ptr(hi), _ptr(comp), _ptr(val)))

    def multigrid(self):
        """collective: builds the level hierarchy of this rank"""
        from ._lib import check, lib
        from .voxelfem_mg import MultigridSolver
        mg = MultigridSolver.__new__(MultigridSolver)
        mg._tps, mg.N, mg._flags = self.tps, self.tps.N, [True, False]
        mg._h = C.c_void_p()
        check(lib().vfem_mg_create_slab(self.tps._h, int(self.levels), self._comm, C.byref(mg._h)))
        mg.lastIterations, mg.lastResidualNorm = 0, 0.0
        self.mg = mg
        return mg

    def dot(self, a, b):
        """global dot product of two local node fields (owned dofs of every rank)"""
        from ._lib import check, lib
        from .voxelfem import _Arg
        n = self.tps.numNodes() * self.tps.N
        aa, bb = _Arg(a, n), _Arg(b, n)
        if aa.ms != bb.ms:
            raise RuntimeError("dot: both arrays must live in the same memory space")
        r = C.c_double()
        check(lib().vfem_mg_dot(self.mg._h, aa.ptr, bb.ptr, aa.ms, C.byref(r)))
        return r.value


def make_nccl_comm(dist, rank, world):
    """NCCL communicator of the C++ solver: rank 0 draws the unique id, torch.distributed broadcasts it."""
    import torch
    from ._lib import check, lib
    buf = np.zeros(128, np.uint8)
    if rank == 0:
        check(lib().vfem_comm_nccl_unique_id(buf.ctypes.data_as(C.c_void_p)))
    t = torch.from_numpy(buf).cuda()
    dist.broadcast(t, 0)
    buf = t.cpu().numpy().copy()
    h = C.c_void_p()
    check(lib().vfem_comm_create_nccl(buf.ctypes.data_as(C.c_void_p), rank, world, C.byref(h)))
    return h


def make_thread_comms(world):
    """`world` ranks as host threads on one device (single-GPU tests of the slab algorithm)."""
    from ._lib import check, lib
    arr = (C.c_void_p * world)()
    check(lib().vfem_comm_create_threads(world, C.cast(arr, C.c_void_p)))
    return [C.c_void_p(arr[r]) for r in range(world)]


class ThreadHostSum:
    """host_sum for ranks that are threads of one process"""

    def __init__(self, world):
        import threading
        self._b = threading.Barrier(world)
        self._vals = [0] * world

    def bind(self, rank):
        def host_sum(v):
            self._vals[rank] = v
            self._b.wait()
            s = sum(self._vals)
            self._b.wait()
            return s
        return host_sum


def dist_host_sum(dist, device):
    import torch

    def host_sum(v):
        t = torch.tensor([v], dtype=torch.int64, device=device)
        dist.all_reduce(t)
        return int(t.item())
    return host_sum
