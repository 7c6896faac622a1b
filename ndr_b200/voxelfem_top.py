"""Filters, constraints, TopologyOptimizationProblem and OCOptimizer of the `pyVoxelFEM` mirror
(VoxelFEM/python_bindings/VoxelFEM.cc:156-183,210-215,258-288)."""
import ctypes as C

import numpy as np

from ._lib import FILTER_CB, check, lib
from .voxelfem import DEVICE, HOST, _Arg, _dev_ptr, _is_device, _np, _ptr
from .benchmark import timed as _timed

_SMOOTHING, _PROJECTION, _CALLBACK = 0, 1, 2


class _Filter:
    _type, _param = None, 0.0
    _dims = None
    _owner = None  # (problem, index) once attached, so attribute updates reach the device problem

    def _push(self):
        if self._owner is not None:
            top, i = self._owner
            check(lib().vfem_top_set_filter_param(top._h, i, float(self._param)))


class SmoothingFilter(_Filter):
    """Box filter of integer radius (TopologyOptimizationFilter.hh:105-162); default radius 1."""
    _type = _SMOOTHING

    def __init__(self):
        self._param = 1.0

    radius = property(lambda s: int(s._param), lambda s, v: (setattr(s, "_param", float(int(v))), s._push())[0])


class ProjectionFilter(_Filter):
    """tanh projection (TopologyOptimizationFilter.hh:55-79); default beta 1."""
    _type = _PROJECTION

    def __init__(self):
        self._param = 1.0

    beta = property(lambda s: s._param, lambda s, v: (setattr(s, "_param", float(v)), s._push())[0])


class PythonFilter(_Filter):
    """Filter defined by Python callbacks apply_cb(in, out) / backprop_cb(in, vars, out) (Filter.hh:81-103)."""
    _type = _CALLBACK

    def __init__(self):
        self.apply_cb = None
        self.backprop_cb = None
        self._error = None  # an exception raised inside the ctypes callback (ctypes would only print it)

        def _cb(ctx, backprop, pin, pvars, pout, n):
            as_np = lambda p: np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(n,))
            try:
                if backprop:
                    if self.backprop_cb is None:
                        raise RuntimeError("Backprop callback must be configured")   # Filter.hh:96
                    self.backprop_cb(as_np(pin), as_np(pvars), as_np(pout))
                else:
                    if self.apply_cb is None:
                        raise RuntimeError("Apply callback must be configured")      # Filter.hh:89
                    self.apply_cb(as_np(pin), as_np(pout))
            except BaseException as e:  # noqa: BLE001
                if self._error is None:
                    self._error = e
                as_np(pout)[:] = np.nan  # never hand stale memory on as a filter result
        self._c_cb = FILTER_CB(_cb)


class LangelaarFilter(PythonFilter):
    """Overhang (additive-manufacturing) filter of Langelaar (TopologyOptimizationFilter.hh:164-278): layers along the LAST
    grid axis, layer 0 on the baseplate; a voxel's density is the smooth minimum of its blueprint density and the smooth
    maximum (P-norm, P = 40, Q = P - 1.58) of its supporting region (voxel below + in-layer neighbours of that voxel).
    Sequential over layers and outside every shipped run: evaluated on the host (numpy, vectorised per layer) through the
    callback-filter slot of the device problem.  The reference's back-propagation takes its layer count from gridDims[1]
    (Filter.hh:216-218) where apply uses gridDims[N-1]; that quirk is mirrored (identical in 2-D)."""
    _eps, _P, _Q = 1e-4, 40.0, 40.0 - 1.58

    def __init__(self):
        super().__init__()
        self.apply_cb, self.backprop_cb = self._apply, self._backprop
        self._smax = self._filtered = None

    def _support(self, a):
        """values of the supporting voxels of every voxel of the next layer: list of (array, validity mask or None)"""
        out = [(a, None)]
        for d in range(a.ndim):
            for sh in (-1, 1):
                r = np.roll(a, -sh, axis=d)
                m = np.ones(a.shape, bool)
                idx = [slice(None)] * a.ndim
                idx[d] = 0 if sh == -1 else -1
                m[tuple(idx)] = False
                out.append((r, m))
        return out

    def _psum(self, layer):
        tot = np.zeros(layer.shape)
        for v, m in self._support(layer):
            tot += np.power(v, self._P) if m is None else np.where(m, np.power(v, self._P), 0.0)
        return tot

    def _smin(self, x1, x2):
        return 0.5 * (x1 + x2 - np.sqrt((x1 - x2) ** 2 + self._eps) + np.sqrt(self._eps))

    def _apply(self, xin, xout):
        if self._dims is None:
            raise RuntimeError("Grid dimensions must be set before filter can be used")
        dims = tuple(self._dims)
        x = xin.reshape(dims)
        out = np.empty(dims)
        smax = np.ones(dims)
        out[..., 0] = x[..., 0]
        for L in range(1, dims[-1]):
            smax[..., L] = np.power(self._psum(out[..., L - 1]), 1.0 / self._Q)
            out[..., L] = self._smin(x[..., L], smax[..., L])
        self._smax, self._filtered = smax, out.copy()
        xout[:] = out.ravel()

    def _backprop(self, din, vars_, dout):
        dims = tuple(self._dims)
        if self._smax is None:
            raise RuntimeError("LangelaarFilter.backprop needs a preceding apply")
        if len(dims) > 2 and dims[1] > dims[-1]:
            raise RuntimeError("NDVector has no layer %d." % (dims[1] - 1))   # the reference's gridDims[1] quirk (Filter.hh:216)
        nl = dims[1]
        d, v, f, sm = din.reshape(dims), vars_.reshape(dims), self._filtered, self._smax
        lam = np.zeros(dims)
        ds2 = lambda x1, x2: 0.5 * (1 + (x1 - x2) / np.sqrt((x1 - x2) ** 2 + self._eps))
        ds1 = lambda x1, x2: 0.5 * (1 - (x1 - x2) / np.sqrt((x1 - x2) ** 2 + self._eps))
        for L in range(nl - 1, -1, -1):
            lam[..., L] = d[..., L]
            if L < nl - 1:
                A = lam[..., L + 1] * ds2(v[..., L + 1], sm[..., L + 1]) * (self._P / self._Q) * np.power(self._psum(f[..., L]), 1.0 / self._Q - 1.0)
                fp = np.power(f[..., L], self._P - 1.0)
                lam[..., L] += A * fp                                   # the voxel below
                for ax in range(A.ndim):                                # its in-layer neighbours
                    for sh in (-1, 1):
                        src = [slice(None)] * A.ndim
                        dst = [slice(None)] * A.ndim
                        src[ax] = slice(1, None) if sh == -1 else slice(None, -1)
                        dst[ax] = slice(None, -1) if sh == -1 else slice(1, None)
                        lam[..., L][tuple(dst)] += A[tuple(src)] * fp[tuple(dst)]
        res = np.zeros(dims)
        res[..., :nl] = (lam * ds1(v, sm))[..., :nl]
        dout[:] = res.ravel()


class TotalVolumeConstraint:
    """c(x) = 1 - mean(xPhys)/volumeFraction >= 0 (TopologyOptimizationConstraint.hh:21-35)."""

    def __init__(self, volumeFraction):
        self.volumeFraction = float(volumeFraction)


def applyFilter(filter, x):
    """pyVoxelFEM.applyFilter(filter, x); the filter must know its grid (set by a TopologyOptimizationProblem)."""
    if filter._dims is None:
        raise RuntimeError("Grid dimensions must be set before filter can be used")  # Filter::checkGridDimensionsAreSet
    if filter._type == _CALLBACK:
        out = np.empty(len(x))
        filter.apply_cb(_np(x), out)
        return out
    dims = _np(filter._dims, np.uint64)
    x = _np(x)
    out = np.empty_like(x)
    check(lib().vfem_apply_filter(filter._type, float(filter._param), len(dims), _ptr(dims), _ptr(x), _ptr(out)))
    return out


class TopologyOptimizationProblem:
    """TopologyOptimizationProblem(simulator, objective, constraints, filters) (VoxelFEM.cc:156-159, Problem.hh:17-204)."""

    def __init__(self, simulator, objective, constraints, filters):
        self._sim, self._objective = simulator, objective
        self._constraints, self._filters = list(constraints), list(filters)
        for c in self._constraints:
            if not isinstance(c, TotalVolumeConstraint):
                raise RuntimeError("only TotalVolumeConstraint is implemented")
        vf = _np([c.volumeFraction for c in self._constraints])
        ft = _np([f._type for f in self._filters], np.int32)
        fp = _np([f._param for f in self._filters])
        nf = len(self._filters)
        cbs = (FILTER_CB * max(nf, 1))(*[(f._c_cb if f._type == _CALLBACK else C.cast(None, FILTER_CB)) for f in self._filters])
        self._h = C.c_void_p()
        check(lib().vfem_top_create(simulator._h, objective._h, len(vf), _ptr(vf), nf, _ptr(ft), _ptr(fp),
                                    C.cast(cbs, C.c_void_p), None, C.byref(self._h)))
        self._cbs = cbs
        for i, f in enumerate(self._filters):
            f._dims = simulator.NbElementsPerDimension()
            f._owner = (self, i)

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:
            try:
                lib().vfem_top_destroy(self._h)
            except Exception:  # interpreter shutdown
                pass
            self._h = None

    objective = property(lambda s: s._objective)
    filters = property(lambda s: s._filters)
    constraints = property(lambda s: s._constraints)

    def numVars(self):
        return lib().vfem_top_num_vars(self._h)

    def _reraise(self):
        """exceptions raised inside PythonFilter callbacks during the last C call (the reference throws through pybind)"""
        for f in self._filters:
            e = getattr(f, "_error", None)
            if e is not None:
                f._error = None
                raise e

    @_timed("setVars")
    def setVars(self, x, forceUpdate=False):
        a = _Arg(x, self.numVars())
        ch = C.c_int(0)
        rc = lib().vfem_top_set_vars(self._h, a.ptr, a.ms, int(forceUpdate), C.byref(ch))
        self._reraise()
        check(rc)
        return bool(ch.value)

    def _get(self, which, out=None):
        if out is not None and _is_device(out):
            check(lib().vfem_top_get_vars(self._h, which, _dev_ptr(out, self.numVars()), DEVICE))
            return out
        v = np.empty(self.numVars())
        check(lib().vfem_top_get_vars(self._h, which, _ptr(v), HOST))
        return v

    def getVars(self, out=None):
        return self._get(0, out)

    def getDensities(self, out=None):
        return self._get(-1, out)

    @_timed("evaluateObjective")
    def evaluateObjective(self):
        J = C.c_double()
        check(lib().vfem_top_objective(self._h, C.byref(J)))
        return J.value

    @_timed("evaluateObjectiveGradient")
    def evaluateObjectiveGradient(self, out=None):
        """numpy array by default; pass a float64 CUDA array as `out` to receive the gradient with no copy."""
        if out is not None and _is_device(out):
            rc = lib().vfem_top_objective_gradient(self._h, _dev_ptr(out, self.numVars()), DEVICE)
            self._reraise(); check(rc)
            return out
        g = np.empty(self.numVars())
        rc = lib().vfem_top_objective_gradient(self._h, _ptr(g), HOST)
        self._reraise(); check(rc)
        return g

    @_timed("evaluateConstraints")
    def evaluateConstraints(self):
        c = np.empty(len(self._constraints))
        check(lib().vfem_top_constraints(self._h, _ptr(c)))
        return c

    @_timed("evaluateConstraintsJacobian")
    def evaluateConstraintsJacobian(self):
        j = np.empty((len(self._constraints), self.numVars()))
        rc = lib().vfem_top_constraints_jacobian(self._h, _ptr(j), HOST)
        self._reraise(); check(rc)
        return j


class OCOptimizer:
    """OCOptimizer(problem).step(m=0.2, ctol=1e-6) (OptimalityCriterion.hh:32-81)."""

    def __init__(self, problem):
        self._p = problem
        check(lib().vfem_oc_reset(problem._h))
        self.lastLambda, self.lastEvaluations = 0.0, 0

    @_timed("OC step")
    def step(self, m=0.2, ctol=1e-6):
        lam, nev = C.c_double(), C.c_int()
        rc = lib().vfem_oc_step(self._p._h, float(m), float(ctol), C.byref(lam), C.byref(nev))
        self._p._reraise(); check(rc)
        self.lastLambda, self.lastEvaluations = lam.value, nev.value
        # the reference prints "objective, constraint, lambda estimate" here (OptimalityCriterion.hh:74)
        print("objective, constraint, lambda estimate: %g\t%g\t%g" % (self._p.evaluateObjective(), self._p.evaluateConstraints()[0], lam.value))
