"""MultigridSolver / objectives / TopologyOptimizationProblem / OCOptimizer / filters / constraints of the
`pyVoxelFEM` mirror (VoxelFEM/python_bindings/VoxelFEM.cc:94-131,156-215,258-288)."""
import ctypes as C

import numpy as np

from . import voxelfem as _v
from ._lib import FILTER_CB, PCG_CB, RESIDUAL_CB, SOLVE_CB, check, lib
from .voxelfem import DEVICE, HOST, _Arg, _dev_ptr, _is_device, _np, _ptr
from .benchmark import timed as _timed


class _LevelView:
    """getSimulator(l): read-only view of a level's grid and Dirichlet mask."""

    def __init__(self, mg, l):
        self._mg, self._l, self.N = mg, l, mg.N

    def numNodes(self):
        return lib().vfem_mg_level_nodes(self._mg._h, self._l)

    def NbElementsPerDimension(self):
        el = np.zeros(self.N, np.uint64)
        check(lib().vfem_mg_level_grid(self._mg._h, self._l, _ptr(el)))
        return [int(e) for e in el]

    def numElements(self):
        return int(np.prod(self.NbElementsPerDimension()))

    @property
    def dirichletMask(self):
        m = np.empty((self.numNodes(), self.N), np.uint8)
        check(lib().vfem_mg_level_dirichlet_mask(self._mg._h, self._l, _ptr(m)))
        return m.astype(bool)


class MultigridSolver:
    def __init__(self, tps, numCoarseningLevels):
        self._tps, self.N = tps, tps.N  # keeps the simulator alive (the reference holds a shared_ptr)
        self._h = C.c_void_p()
        check(lib().vfem_mg_create(tps._h, int(numCoarseningLevels), C.byref(self._h)))
        self._flags = [True, False]

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:
            try:
                lib().vfem_mg_destroy(self._h)
            except Exception:  # interpreter shutdown
                pass
            self._h = None

    def numLevels(self):
        return lib().vfem_mg_num_levels(self._h)

    def getSimulator(self, l):
        return self._tps if l == 0 else _LevelView(self, l)

    def _n(self, l):
        return lib().vfem_mg_level_nodes(self._h, l) * self.N

    def _shape(self, l):
        return (lib().vfem_mg_level_nodes(self._h, l), self.N)

    def computeResidual(self, l, u, b):
        au, ab = _Arg(u, self._n(l)), _Arg(b, self._n(l))
        out = np.empty(self._shape(l))
        check(lib().vfem_mg_compute_residual(self._h, l, au.ptr, ab.ptr, _ptr(out), HOST))
        return out

    def applyK(self, l, u):
        au = _Arg(u, self._n(l))
        out = np.empty(self._shape(l))
        check(lib().vfem_mg_apply_k(self._h, l, au.ptr, _ptr(out), HOST))
        return out

    def zeroOutDirichletComponents(self, l, u):
        out = _np(u).copy().reshape(self._shape(l))
        check(lib().vfem_mg_zero_dirichlet(self._h, l, _ptr(out), HOST))
        return out

    def smoothingMulticoloredGS(self, l, u, b, forward=True):
        """One multicoloured Gauss-Seidel sweep (MG.hh:336-340), the smoother used inside the cycles."""
        out = _np(u).copy().reshape(self._shape(l))
        ab = _Arg(b, self._n(l))
        check(lib().vfem_mg_smooth_colored(self._h, l, _ptr(out), ab.ptr, int(forward), HOST))
        return out

    def restriction(self, l, v):
        av = _Arg(v, self._n(l))
        out = np.empty(self._shape(l + 1))
        check(lib().vfem_mg_restrict(self._h, l, av.ptr, _ptr(out), HOST))
        return out

    def interpolation(self, l, xc, accumulate_into=None):
        ac = _Arg(xc, self._n(l + 1))
        out = np.empty(self._shape(l)) if accumulate_into is None else _np(accumulate_into).copy().reshape(self._shape(l))
        check(lib().vfem_mg_interpolate(self._h, l, ac.ptr, _ptr(out), int(accumulate_into is not None), HOST))
        return out

    def updateElementStiffnessMatrices(self):
        check(lib().vfem_mg_update_stiffness(self._h))

    updateBlockKs = updateElementStiffnessMatrices

    def setSymmetricGaussSeidel(self, symmetric):
        check(lib().vfem_mg_set_symmetric_gs(self._h, int(symmetric)))

    def _set_flag(self, i, v):
        self._flags[i] = bool(v)
        check(lib().vfem_mg_set_flags(self._h, int(self._flags[0]), int(self._flags[1])))

    buildBlockStiffnessMatrices = property(lambda s: s._flags[0], lambda s, v: s._set_flag(0, v))
    buildFinestBlockStiffnessMatrix = property(lambda s: s._flags[1], lambda s, v: s._set_flag(1, v))

    def levelStencil(self, l):
        nst = 3 ** self.N
        out = np.empty((nst, self.N, self.N, lib().vfem_mg_level_nodes(self._h, l)))
        check(lib().vfem_mg_level_stencil(self._h, l, _ptr(out)))
        return out

    def debug_get_x(self, l):
        x = np.empty(self._shape(l))
        check(lib().vfem_mg_debug_get(self._h, l, _ptr(x), None))
        return x

    def debug_get_b(self, l):
        b = np.empty(self._shape(l))
        check(lib().vfem_mg_debug_get(self._h, l, None, _ptr(b)))
        return b

    @_timed("MG Solver")
    def solve(self, u, f, numSteps, numSmoothingSteps, stiffnessUpdated=False, zeroDirichlet=False, it_callback=None,
              fullMultigrid=False):
        n, shape = self._n(0), self._shape(0)
        au, af = _Arg(u, n), _Arg(f, n)
        out = np.empty(shape)
        cb = SOLVE_CB(lambda ctx, it, xp: it_callback(it, np.ctypeslib.as_array(C.cast(xp, C.POINTER(C.c_double)), shape=shape).copy())) \
            if it_callback else C.cast(None, SOLVE_CB)
        check(lib().vfem_mg_solve(self._h, au.ptr, af.ptr, int(numSteps), int(numSmoothingSteps), int(stiffnessUpdated), int(zeroDirichlet),
                                  cb, None, int(fullMultigrid), _ptr(out), HOST))
        return out

    @_timed("CG Iterations")
    def preconditionedConjugateGradient(self, u, b, maxIter, tol, it_callback=None, mgIterations=1, mgSmoothingIterations=1,
                                        fullMultigrid=False, out=None, return_info=False):
        n, shape = self._n(0), self._shape(0)
        au, ab = _Arg(u, n), _Arg(b, n)
        if au.ms != ab.ms:
            raise RuntimeError("u and b must live in the same memory space")
        if au.ms == DEVICE:
            if out is None:
                raise RuntimeError("device-resident solve needs a preallocated `out` device array")
            optr = _dev_ptr(out, n)
        else:
            out = np.empty(shape)
            optr = _ptr(out)

        def _cb(ctx, it, xp, rp, rn):
            as_np = lambda p: np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=shape).copy()
            it_callback(it, as_np(xp), as_np(rp))
        cb = PCG_CB(_cb) if it_callback else C.cast(None, PCG_CB)
        its, rn = C.c_size_t(0), C.c_double(0)
        check(lib().vfem_mg_pcg(self._h, au.ptr, ab.ptr, int(maxIter), float(tol), cb, None, 1, int(mgIterations), int(mgSmoothingIterations),
                                int(fullMultigrid), optr, au.ms, C.byref(its), C.byref(rn)))
        self.lastIterations, self.lastResidualNorm = its.value, rn.value
        return (out, its.value, rn.value) if return_info else out

    def smoothing(self, l, u, b):
        """Serial lexicographic Gauss-Seidel on a copy of u, after updateElementStiffnessMatrices (VoxelFEM.cc:99-104,
        MG.hh:269-282) - a debug helper of the reference; the cycles use smoothingMulticoloredGS."""
        self.updateElementStiffnessMatrices()
        out = _np(u).copy().reshape(self._shape(l))
        ab = _Arg(b, self._n(l))
        check(lib().vfem_mg_smooth_serial(self._h, l, _ptr(out), ab.ptr, 1, HOST))
        return out

    def debugMulticolorVisit(self):
        """Visit order of the multicoloured sweep (MG.hh:328-334): colours in row-major local-node order."""
        nper = [e + 1 for e in self._tps.NbElementsPerDimension()]
        idx = np.indices(nper).reshape(self.N, -1)
        colour = np.zeros(idx.shape[1], np.int64)
        for d in range(self.N):
            colour = colour * 2 + (idx[d] & 1)
        order = np.empty(idx.shape[1], np.int64)
        order[np.argsort(colour, kind="stable")] = np.arange(idx.shape[1])
        return order


class MultigridComplianceObjective:
    """MultigridComplianceObjective(mg_solver) (VoxelFEM.cc:164-167, Objective.hh:70-102)."""
    _FIELDS = ("cgIter", "tol", "mgIterations", "mgSmoothingIterations", "fullMultigrid", "zeroInit")

    def __init__(self, mg_solver):
        object.__setattr__(self, "mg", mg_solver)
        object.__setattr__(self, "_h", C.c_void_p())
        object.__setattr__(self, "_rcb", None)
        object.__setattr__(self, "_rcb_c", None)
        check(lib().vfem_obj_create(mg_solver._h, C.byref(self._h)))

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:
            try:
                lib().vfem_obj_destroy(self._h)
            except Exception:  # interpreter shutdown
                pass
            object.__setattr__(self, "_h", None)

    def _params(self):
        a, b, c, d, e, f = C.c_size_t(), C.c_double(), C.c_size_t(), C.c_size_t(), C.c_int(), C.c_int()
        check(lib().vfem_obj_get_params(self._h, C.byref(a), C.byref(b), C.byref(c), C.byref(d), C.byref(e), C.byref(f)))
        return [a.value, b.value, c.value, d.value, bool(e.value), bool(f.value)]

    def __getattr__(self, name):
        if name in MultigridComplianceObjective._FIELDS:
            return self._params()[self._FIELDS.index(name)]
        if name == "residual_cb":
            return self._rcb
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name in self._FIELDS:
            p = self._params()
            p[self._FIELDS.index(name)] = value
            check(lib().vfem_obj_set_params(self._h, int(p[0]), float(p[1]), int(p[2]), int(p[3]), int(p[4]), int(p[5])))
        elif name == "residual_cb":
            object.__setattr__(self, "_rcb", value)
            cb = RESIDUAL_CB(lambda ctx, it, r: value(it, r)) if value else C.cast(None, RESIDUAL_CB)
            object.__setattr__(self, "_rcb_c", cb)
            check(lib().vfem_obj_set_residual_cb(self._h, cb, None))
        else:
            raise AttributeError("can't set attribute " + name)

    def compliance(self):
        J = C.c_double()
        check(lib().vfem_obj_compliance(self._h, C.byref(J)))
        return J.value

    def _sim(self):
        return self.mg._tps if self.mg is not None else self._tps

    def gradient(self, out=None):
        ne = self._sim().numElements()
        if out is not None and _is_device(out):
            check(lib().vfem_obj_gradient(self._h, _dev_ptr(out, ne), DEVICE))
            return out
        g = np.empty(ne)
        check(lib().vfem_obj_gradient(self._h, _ptr(g), HOST))
        return g

    def u(self):
        u = np.empty((self._sim().numNodes(), self._sim().N))
        check(lib().vfem_obj_get_u(self._h, _ptr(u), HOST))
        return u

    def f(self):
        f = np.empty((self._sim().numNodes(), self._sim().N))
        check(lib().vfem_obj_get_f(self._h, _ptr(f), HOST))
        return f

    def updateCache(self, xPhys):
        a = _Arg(xPhys, self._sim().numElements())
        check(lib().vfem_obj_update_cache(self._h, a.ptr, a.ms))

    def lastSolve(self):
        its, rn = C.c_size_t(), C.c_double()
        check(lib().vfem_obj_last_solve(self._h, C.byref(its), C.byref(rn)))
        return its.value, rn.value


class ComplianceObjective(MultigridComplianceObjective):
    """pyVoxelFEM.ComplianceObjective(simulator) (VoxelFEM.cc:160-163, Objective.hh:25-68): every updateCache is a direct
    solve (block LDL^T on the device where the reference calls CHOLMOD) - the `--mgl 0` path of fem.py:44-45."""
    _FIELDS = ()

    def __init__(self, simulator):
        object.__setattr__(self, "mg", None)
        object.__setattr__(self, "_tps", simulator)
        object.__setattr__(self, "_h", C.c_void_p())
        object.__setattr__(self, "_rcb", None)
        object.__setattr__(self, "_rcb_c", None)
        check(lib().vfem_obj_create_direct(simulator._h, C.byref(self._h)))
