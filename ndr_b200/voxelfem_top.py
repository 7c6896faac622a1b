"""Filters, constraints, TopologyOptimizationProblem and OCOptimizer of the `pyVoxelFEM` mirror
(VoxelFEM/python_bindings/VoxelFEM.cc:156-183,210-215,258-288)."""
import ctypes as C

import numpy as np

from ._lib import FILTER_CB, check, lib
from .voxelfem import DEVICE, HOST, _Arg, _dev_ptr, _is_device, _np, _ptr

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


class LangelaarFilter(_Filter):
    def __init__(self):
        raise RuntimeError("LangelaarFilter is outside the device hot path (SURVEY 8f row 4)")


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

    def evaluateObjective(self):
        J = C.c_double()
        check(lib().vfem_top_objective(self._h, C.byref(J)))
        return J.value

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

    def evaluateConstraints(self):
        c = np.empty(len(self._constraints))
        check(lib().vfem_top_constraints(self._h, _ptr(c)))
        return c

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

    def step(self, m=0.2, ctol=1e-6):
        lam, nev = C.c_double(), C.c_int()
        rc = lib().vfem_oc_step(self._p._h, float(m), float(ctol), C.byref(lam), C.byref(nev))
        self._p._reraise(); check(rc)
        self.lastLambda, self.lastEvaluations = lam.value, nev.value
        # the reference prints "objective, constraint, lambda estimate" here (OptimalityCriterion.hh:74)
        print("objective, constraint, lambda estimate: %g\t%g\t%g" % (self._p.evaluateObjective(), self._p.evaluateConstraints()[0], lam.value))
