"""GPU parity at the BASELINE.json configurations themselves.

 * config 2 (3-D cantilever 128x64x64): K-apply / PCG against the oracle on the full grid;
 * config 1 (2-D MBB 192x64): the whole chain against the oracle;
 * the reference's own 3-D MG-PCG logs (logs/slurm/gt/c1001.log, b1000.log -> tests/golden/oc_trajectories.json):
   the GPU runs the exact call order of fem.py:31-87 and must print the reference's compliance values - no oracle
   in between;
 * 50 optimality-criterion iterations GPU vs oracle (max-abs density <= 1e-8 at every step).
Everything goes through the C ABI (ctypes mirror of pyVoxelFEM)."""
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, MATERIAL, bc_path

pytestmark = pytest.mark.gpu

TRAJ = json.load(open(os.path.join(GOLDEN, "oc_trajectories.json")))


def rel(a, b):
    return float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(b)), 1e-300))


def pair(O, dom, grid, bc, rho):
    import ndr_b200.pyVoxelFEM as V
    dim = len(grid)
    o = O.OracleSim([0] * dim, dom, grid)
    o.read_material(MATERIAL)
    g = V.TensorProductSimulator([1] * dim, [np.zeros(dim), np.array(dom, float)], grid)
    g.readMaterial(MATERIAL)
    o.set_densities(rho); g.setElementDensities(rho)
    o.apply_bc_file(bc_path(bc)); g.applyDisplacementsAndLoadsFromFile(bc_path(bc))
    o.set_simp(1.0, 1e-4, 3.0)
    g.E_0, g.E_min, g.gamma = 1.0, 1e-4, 3.0
    return o, g


def test_config2_cantilever_128x64x64(oracle_mod):
    """BASELINE configs[1]: K-apply <= 1e-13, production PCG (tol 1e-4) with equal iteration count and residual,
    converged PCG (tol 1e-12) with displacement and compliance <= 1e-10 (north_star)."""
    grid = [128, 64, 64]
    rho = np.random.default_rng(88).uniform(0, 1, int(np.prod(grid)))
    o, g = pair(oracle_mod, [2, 1, 1], grid, "3d/cantilever_flexion.bc", rho)
    assert g.applyInfo()["march"] == 1  # the marching kernel is the one under test here
    u = np.random.default_rng(89).standard_normal((o.nn, 3))
    assert rel(g.applyK(u), o.apply_k(u)) < 1e-13
    assert rel(g.complianceGradient(u), o.compliance_gradient(u)) < 1e-13
    o.mg_create(3)
    mg = g.multigridSolver(3)
    f = o.load_vector()
    np.testing.assert_array_equal(g.buildLoadVector(), f)
    u0 = np.zeros_like(f)
    kw = dict(mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True, return_info=True)
    got, kg, rn = mg.preconditionedConjugateGradient(u0, f, 100, 1e-4, **kw)
    ref, ko, hist = o.mg_pcg(u0, f, 100, 1e-4, 1, 2, True)
    assert kg == ko and abs(rn - hist[-1]) <= 1e-6 * hist[-1]
    assert rel(got, ref) < 1e-9
    got, kg, _ = mg.preconditionedConjugateGradient(u0, f, 100, 1e-12, **kw)
    ref, ko, _ = o.mg_pcg(u0, f, 100, 1e-12, 1, 2, True)
    assert rel(got, ref) < 1e-10
    assert abs(np.sum(f * got) - np.sum(f * ref)) <= 1e-10 * abs(np.sum(f * ref))
    assert rel(g.complianceGradient(got), o.compliance_gradient(ref)) < 1e-9


def test_config1_mbb_192x64(oracle_mod):
    """BASELINE configs[0]: 2-D MBB beam 192x64, SIMP p = 3, MG-PCG + optimality criterion, whole chain vs the oracle."""
    import ndr_b200.pyVoxelFEM as V
    grid, vol, levels = [192, 64], 0.3, 3
    o, g = pair(oracle_mod, [3, 1], grid, "2d/mbb_beam.bc", np.full(192 * 64, vol))
    u = np.random.default_rng(89).standard_normal((o.nn, 2))
    assert rel(g.applyK(u), o.apply_k(u)) < 1e-13
    o.mg_create(levels)
    mg = g.multigridSolver(levels)
    f = o.load_vector()
    u0 = np.zeros_like(f)
    got, kg, _ = mg.preconditionedConjugateGradient(u0, f, 100, 1e-12, mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True, return_info=True)
    ref, ko, _ = o.mg_pcg(u0, f, 100, 1e-12, 1, 2, True)
    assert rel(got, ref) < 1e-10
    o.top_create(True, vol, [("smoothing", 1), ("projection", 1.0)])
    o.top_set_solver(100, 1e-11, 1, 2, True, False)
    obj = V.MultigridComplianceObjective(mg)
    top = V.TopologyOptimizationProblem(g, obj, [V.TotalVolumeConstraint(vol)], [V.SmoothingFilter(), V.ProjectionFilter()])
    obj.tol = 1e-11
    x0 = np.full(o.ne, vol)
    o.top_set_vars(x0); top.setVars(x0)
    oc = V.OCOptimizer(top)
    for step in range(10):
        lam_o, _ = o.oc_step()
        oc.step()
        assert abs(oc.lastLambda - lam_o) <= 1e-9 * lam_o, step
        assert np.max(np.abs(top.getVars() - o.top_vars(0))) <= 1e-8, step
        Jo = o.top_objective()
        assert abs(top.evaluateObjective() - Jo) <= 1e-9 * Jo, step


def fem_py_run(V, cfg, mgl, nsteps):
    """The statements of helpers/ipopt_helpers.py:7-15 and fem.py:31-87 (optimizer == 'OC'), in their order."""
    vol = cfg["maxVolume"][0]
    constraints = [V.TotalVolumeConstraint(vol)]
    filters = [V.SmoothingFilter(), V.ProjectionFilter()]
    tps = V.TensorProductSimulator(cfg["orderFEM"], cfg["domainCorners"], cfg["gridDimensions"])
    tps.readMaterial(MATERIAL)
    tps.setUniformDensities(vol)
    tps.applyDisplacementsAndLoadsFromFile(bc_path(cfg["BC_PATH"][len("bcs/"):]))
    tps.E_0, tps.E_min, tps.gamma = cfg["E0"], cfg["Emin"], cfg["SIMPExponent"]
    objective = V.MultigridComplianceObjective(tps.multigridSolver(mgl))   # solves once at the default tol 1e-5
    top = V.TopologyOptimizationProblem(tps, objective, constraints, filters)
    objective.tol = 1e-4
    objective.mgIterations = 1
    objective.fullMultigrid = True
    objective.zeroInit = False
    objective.mgSmoothingIterations = 2
    oco = V.OCOptimizer(top)
    top.setVars(tps.getDensities())
    comp = []
    for idx in range(nsteps + 1):
        comp.append(2.0 * top.evaluateObjective())
        if idx < nsteps:
            oco.step()
    return comp, top


@pytest.mark.parametrize("log,nsteps", [("c1001", 2), ("b1000", 1)])
def test_reference_3d_log_trajectory(log, nsteps):
    """logs/slurm/gt/c1001.log:137-187 (256x128x128 cantilever, V = 0.5, mgl 3) and b1000.log:141-191 (320x160x80
    bridge, V = 0.4): the reference's printed `Compliance loss` of steps 0..n.  Its solves stop at tol 1e-4, so step 0
    agrees to ~1e-4 relative and later steps (which inherit the design differences) to ~1e-3."""
    import ndr_b200.pyVoxelFEM as V
    gold = TRAJ[log]
    cfg = eval(gold["config"])
    comp, _ = fem_py_run(V, cfg, gold["multigrid_levels"], nsteps)
    for k, c in enumerate(comp):
        tol = 2e-4 if k == 0 else 2e-3
        assert abs(c - gold["compliance"][k]) <= tol * gold["compliance"][k], (k, c, gold["compliance"][k])


def test_oc_50_iterations_vs_oracle(oracle_mod):
    """Design after N = 50 optimality-criterion iterations on 64x32x32: max-abs density <= 1e-8 at EVERY step
    (north_star); tight solver tolerance on both sides so the iterates are the converged displacement."""
    import ndr_b200.pyVoxelFEM as V
    grid, vol = [64, 32, 32], 0.4
    o, g = pair(oracle_mod, [2, 1, 1], grid, "3d/cantilever_flexion.bc", np.full(int(np.prod(grid)), vol))
    o.mg_create(3)
    o.top_create(True, vol, [("smoothing", 1), ("projection", 1.0)])
    o.top_set_solver(100, 1e-11, 1, 2, True, False)
    obj = V.MultigridComplianceObjective(g.multigridSolver(3))
    top = V.TopologyOptimizationProblem(g, obj, [V.TotalVolumeConstraint(vol)], [V.SmoothingFilter(), V.ProjectionFilter()])
    obj.tol = 1e-11
    x0 = np.full(o.ne, vol)
    o.top_set_vars(x0); top.setVars(x0)
    oc = V.OCOptimizer(top)
    worst = 0.0
    for step in range(50):
        lam_o, _ = o.oc_step()
        oc.step()
        d = float(np.max(np.abs(top.getVars() - o.top_vars(0))))
        worst = max(worst, d)
        assert d <= 1e-8, (step, d)
        Jo = o.top_objective()
        assert abs(top.evaluateObjective() - Jo) <= 1e-8 * Jo, step
    print("50 OC steps: worst max-abs density difference %.3e" % worst)


def test_config2_oc_iterations_vs_oracle(oracle_mod):
    """BASELINE configs[1] at its full size, 128x64x64, with the production solver settings of fem.py (tol 1e-4, FMG, 2 sweeps,
    warm start): the design after each of 8 optimality-criterion iterations against the oracle, max-abs density <= 1e-8
    (north_star) - the two sides run the same PCG iterations, so they agree far below the solver tolerance."""
    import ndr_b200.pyVoxelFEM as V
    grid, vol = [128, 64, 64], 0.4
    o, g = pair(oracle_mod, [2, 1, 1], grid, "3d/cantilever_flexion.bc", np.full(int(np.prod(grid)), vol))
    o.mg_create(3)
    o.top_create(True, vol, [("smoothing", 1), ("projection", 1.0)])
    o.top_set_solver(100, 1e-4, 1, 2, True, False)
    obj = V.MultigridComplianceObjective(g.multigridSolver(3))
    top = V.TopologyOptimizationProblem(g, obj, [V.TotalVolumeConstraint(vol)], [V.SmoothingFilter(), V.ProjectionFilter()])
    obj.tol = 1e-4
    x0 = np.full(o.ne, vol)
    o.top_set_vars(x0); top.setVars(x0)
    oc = V.OCOptimizer(top)
    worst = 0.0
    for step in range(8):
        lam_o, _ = o.oc_step()
        oc.step()
        assert abs(oc.lastLambda - lam_o) <= 1e-9 * lam_o, step
        d = float(np.max(np.abs(top.getVars() - o.top_vars(0))))
        worst = max(worst, d)
        assert d <= 1e-8, (step, d)
        Jo = o.top_objective()
        assert abs(top.evaluateObjective() - Jo) <= 1e-9 * Jo, step
    print("config 2, 8 OC steps: worst max-abs density difference %.3e" % worst)


def test_config2_50_oc_iterations_trajectory():
    """BASELINE configs[1] "50 OC iterations" at 128x64x64, GPU only (fem.py protocol, tol 1e-4): the compliance
    trajectory is compared with the committed one (tests/golden/config2_oc50_gpu.json, written by the first run of this
    test on a B200 - a regression pin, not a reference golden) and must decrease like the reference's c1001 run does."""
    import ndr_b200.pyVoxelFEM as V
    cfg = eval(TRAJ["c1001"]["config"])
    cfg["gridDimensions"] = [128, 64, 64]
    comp, top = fem_py_run(V, cfg, 3, 50)
    assert all(np.isfinite(comp)) and comp[50] < 0.2 * comp[0]
    assert abs(top.evaluateConstraints()[0]) <= 1e-6
    out = os.environ.get("VFEM_WRITE_GOLDEN")
    p = os.path.join(GOLDEN, "config2_oc50_gpu.json")
    if out:
        os.makedirs(os.path.dirname(out), exist_ok=True)
        json.dump({"grid": [128, 64, 64], "compliance": comp}, open(out, "w"))
    if os.path.exists(p):
        ref = json.load(open(p))["compliance"]
        for k in range(51):
            assert abs(comp[k] - ref[k]) <= 5e-3 * ref[k], (k, comp[k], ref[k])
