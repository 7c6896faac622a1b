"""GPU parity of the density pipeline: filters, objective / gradient / constraint chain, OC steps."""
import numpy as np
import pytest

from conftest import MATERIAL, bc_path
from test_gpu_parity import CASES, make_pair, rel

pytestmark = pytest.mark.gpu


def make_problem(O, name, filters=True, tol=1e-10):
    import ndr_b200.pyVoxelFEM as V
    o, g, levels = make_pair(O, name, rho_mode="uniform")
    vol = 0.4
    o.mg_create(levels)
    o.top_create(True, vol, [("smoothing", 1), ("projection", 1.0)] if filters else [])
    o.top_set_solver(100, tol, 1, 2, True, False)
    mg = g.multigridSolver(levels)
    obj = V.MultigridComplianceObjective(mg)
    fl = [V.SmoothingFilter(), V.ProjectionFilter()] if filters else []
    top = V.TopologyOptimizationProblem(g, obj, [V.TotalVolumeConstraint(vol)], fl)
    obj.tol = tol
    assert obj.cgIter == 100 and obj.mgSmoothingIterations == 2 and obj.fullMultigrid and not obj.zeroInit
    return o, g, top, obj


def test_filters_standalone(oracle_mod):
    import ndr_b200.pyVoxelFEM as V
    O = oracle_mod
    rng = np.random.default_rng(3)
    for dims in ([7, 5], [6, 5, 4], [1, 3, 2]):
        x = rng.uniform(0, 1, int(np.prod(dims)))
        for kind, param, F in (("smoothing", 1, V.SmoothingFilter), ("smoothing", 2, V.SmoothingFilter), ("projection", 2.5, V.ProjectionFilter)):
            f = F()
            if kind == "smoothing":
                f.radius = param
            else:
                f.beta = param
            f._dims = dims
            assert rel(V.applyFilter(f, x), O.apply_filter(kind, param, dims, x)) < 1e-14, (dims, kind, param)


@pytest.mark.parametrize("name", ["cant3d", "mbb2d"])
@pytest.mark.parametrize("filters", [True, False])
def test_objective_gradient_constraints(oracle_mod, name, filters):
    o, g, top, obj = make_problem(oracle_mod, name, filters)
    x = np.random.default_rng(5).uniform(0.2, 0.9, o.ne)
    assert o.top_set_vars(x) and top.setVars(x)
    assert not top.setVars(x)  # unchanged variables are skipped (Problem.hh:50-51)
    assert rel(top.getDensities(), o.top_vars(-1)) < 1e-14
    Jo = o.top_objective()
    assert abs(top.evaluateObjective() - Jo) <= 1e-9 * abs(Jo)
    assert rel(obj.u(), o.top_u()) < 1e-8
    assert rel(top.evaluateObjectiveGradient(), o.top_gradient()) < 1e-8
    assert abs(top.evaluateConstraints()[0] - o.top_constraint()) < 1e-13
    assert rel(top.evaluateConstraintsJacobian()[0], o.top_constraint_jac()) < 1e-13


@pytest.mark.parametrize("name", ["cant3d", "mbb2d"])
def test_oc_trajectory(oracle_mod, name):
    """Design after N OC iterations: <= 1e-8 max-abs density difference (north_star), tight solver tolerance so
    both sides converge to the same displacement field."""
    import ndr_b200.pyVoxelFEM as V
    o, g, top, obj = make_problem(oracle_mod, name, True, tol=1e-11)
    x0 = np.full(o.ne, 0.4)
    o.top_set_vars(x0); top.setVars(x0)
    oc = V.OCOptimizer(top)
    for step in range(5):
        lam_o, _ = o.oc_step()
        oc.step()
        assert abs(oc.lastLambda - lam_o) <= 1e-9 * lam_o, step
        assert np.max(np.abs(top.getVars() - o.top_vars(0))) <= 1e-8, step
        Jo = o.top_objective()
        assert abs(top.evaluateObjective() - Jo) <= 1e-8 * Jo, step


def test_python_filter_callback(oracle_mod):
    """PythonFilter round trip (Filter.hh:81-103): identity callback leaves the pipeline unchanged."""
    import ndr_b200.pyVoxelFEM as V
    o, g, levels = make_pair(oracle_mod, "mbb2d", rho_mode="uniform")
    mg = g.multigridSolver(levels)
    obj = V.MultigridComplianceObjective(mg)
    pf = V.PythonFilter()
    calls = []

    def apply_cb(xin, xout):
        calls.append("a"); xout[:] = 0.5 * xin

    def backprop_cb(din, vars_, dout):
        calls.append("b"); dout[:] = 0.5 * din
    pf.apply_cb, pf.backprop_cb = apply_cb, backprop_cb
    top = V.TopologyOptimizationProblem(g, obj, [V.TotalVolumeConstraint(0.4)], [pf])
    x = np.full(g.numElements(), 0.8)
    top.setVars(x)
    np.testing.assert_allclose(top.getDensities(), 0.4)
    g1 = top.evaluateObjectiveGradient()
    assert calls.count("a") == 1 and calls.count("b") == 1 and np.all(g1 < 0)
