"""The drop-in boundary as the reference's scripts see it (SURVEY 8b "Sibling modules the scripts import"):
`import pyVoxelFEM, MeshFEM, mesh, parallelism, benchmark` (+ `cyipopt`) from ndr_b200/shims, the MSH writer / parser,
and - on the GPU - the statements of helpers/ipopt_helpers.py:7-15 and fem.py:31-87 executed through those imports (with the
reference's REAL helper modules when /root/reference is present; they never are on the GPU box, where the flow runs on a
restatement of the same statements)."""
import importlib
import os
import sys
import types

import numpy as np
import pytest

from conftest import MATERIAL, ROOT, bc_path

SHIMS = os.path.join(ROOT, "ndr_b200", "shims")
REF_HELPERS = "/root/reference/VoxelFEM/python/helpers"
SHIM_MODULES = ("pyVoxelFEM", "MeshFEM", "mesh", "parallelism", "benchmark", "cyipopt")


@pytest.fixture
def shim_path():
    saved_path, saved_mods = list(sys.path), {k: sys.modules.get(k) for k in SHIM_MODULES + ("ipopt_helpers", "history_helpers", "mangled_names",
                                                                                            "multiscale_helpers", "matplotlib", "matplotlib.pyplot")}
    for k in saved_mods:
        sys.modules.pop(k, None)
    sys.path.insert(0, SHIMS)
    yield
    sys.path[:] = saved_path
    for k, v in saved_mods.items():
        sys.modules.pop(k, None)
        if v is not None:
            sys.modules[k] = v


def test_sibling_modules_import(shim_path):
    """fem.py:12-14, utils.py:12-13, train_voxelfem.py:20: the import lines of the reference's scripts"""
    import pyVoxelFEM, MeshFEM, mesh, parallelism, benchmark  # noqa: E401,F401
    import cyipopt
    assert os.path.dirname(os.path.abspath(mesh.__file__)) == SHIMS
    for name in ("TensorProductSimulator", "MultigridComplianceObjective", "ComplianceObjective", "TopologyOptimizationProblem", "OCOptimizer",
                 "TotalVolumeConstraint", "SmoothingFilter", "ProjectionFilter", "LangelaarFilter", "PythonFilter", "applyFilter",
                 "benchmark_reset", "benchmark_report"):
        assert hasattr(pyVoxelFEM, name), name
    parallelism.set_max_num_tbb_threads(4)
    parallelism.set_gradient_assembly_num_threads(4)
    assert hasattr(mesh, "MSHFieldWriter") and hasattr(mesh, "MSHFieldParser3")
    nlp = cyipopt.problem(n=4, m=1, problem_obj=None, lb=np.zeros(4), ub=np.ones(4), cl=[0.0], cu=None)
    nlp.addOption("max_iter", 3)


@pytest.mark.parametrize("binary", [True, False])
def test_msh_round_trip(shim_path, tmp_path, binary):
    """MSH 2.2 as MeshFEM writes it (MeshIO.cc:533-612, MSHFieldWriter.hh:128-196): mesh + per-element / per-node fields"""
    import mesh
    rng = np.random.default_rng(1)
    V = rng.uniform(0, 1, (27, 3))
    F = np.array([[0, 1, 4, 3, 9, 10, 13, 12], [1, 2, 5, 4, 10, 11, 14, 13]])
    p = str(tmp_path / "a.msh")
    w = mesh.MSHFieldWriter(p, V, F, binary)
    rho, disp = rng.uniform(0, 1, 2), rng.standard_normal((27, 3))
    w.addField("density", rho)
    w.addField("u", disp, mesh.MSHFieldWriter.DomainType.PER_NODE)
    with pytest.raises(RuntimeError, match="Invalid field domain size"):
        w.addField("bad", np.zeros(5))
    w.close()
    head = open(p, "rb").read(20)
    assert head.startswith(b"$MeshFormat\n2.2 %d 8\n" % (1 if binary else 0))
    r = mesh.MSHFieldParser3(p)
    np.testing.assert_array_equal(r.vertices(), V)
    np.testing.assert_array_equal(r.elements(), F)
    np.testing.assert_array_equal(r.scalarField("density"), rho)
    np.testing.assert_array_equal(r.vectorField("u"), disp)
    with pytest.raises(RuntimeError):
        r.scalarField("nope")


def _reference_helpers():
    """the reference's real ipopt_helpers (imports history_helpers -> matplotlib, which this image lacks: stubbed)"""
    if not os.path.isdir(REF_HELPERS):
        return None
    for name in ("matplotlib", "matplotlib.pyplot"):
        if name not in sys.modules:
            sys.modules[name] = types.ModuleType(name)
    sys.path.insert(1, REF_HELPERS)
    return importlib.import_module("ipopt_helpers")


@pytest.mark.gpu
def test_fem_py_oc_loop_through_the_shims(shim_path, tmp_path):
    """fem.py:31-87 (`ground_truth_topopt`, optimizer == 'OC') statement for statement on 32x16x16, incl. the checkpoint
    writes of utils.save_densities (mesh.MSHFieldWriter(*tps.getMesh()).addField('density', ...)) and a read-back through
    tps.readDensities.  `initializeTensorProductSimulator` / `initializeIpoptProblem` are the reference's own functions when
    /root/reference exists."""
    import pyVoxelFEM
    import mesh
    helpers = _reference_helpers()
    orderFEM, domainCorners, gridDimensions = [1, 1, 1], [[0, 0, 0], [2, 1, 1]], [32, 16, 16]
    maxVolume, E0, Emin, SIMPExponent, multigrid_levels, max_iter = 0.5, 1, 1e-4, 3, 2, 3
    BC_PATH = bc_path("3d/cantilever_flexion.bc")
    constraints = [pyVoxelFEM.TotalVolumeConstraint(maxVolume)]
    filters = [pyVoxelFEM.SmoothingFilter(), pyVoxelFEM.ProjectionFilter()]
    if helpers is not None:
        tps = helpers.initializeTensorProductSimulator(orderFEM, domainCorners, gridDimensions, maxVolume, E0, Emin, SIMPExponent, MATERIAL, BC_PATH)
    else:  # helpers/ipopt_helpers.py:7-15
        tps = pyVoxelFEM.TensorProductSimulator(orderFEM, domainCorners, gridDimensions)
        tps.readMaterial(MATERIAL)
        tps.setUniformDensities(maxVolume)
        tps.applyDisplacementsAndLoadsFromFile(BC_PATH)
        tps.E_0, tps.E_min, tps.gamma = E0, Emin, SIMPExponent
    objective = pyVoxelFEM.MultigridComplianceObjective(tps.multigridSolver(multigrid_levels))
    top = pyVoxelFEM.TopologyOptimizationProblem(tps, objective, constraints, filters)
    if helpers is not None:
        nonLinearProblem, problemObj = helpers.initializeIpoptProblem(top)
        problemObj.beta_interval, problemObj.beta_scaler, problemObj.radius_interval, problemObj.radius_scaler = [1, 1, 1, 1]
    x0 = tps.getDensities()
    objective.tol = 1e-4
    objective.mgIterations = 1
    objective.fullMultigrid = True
    objective.zeroInit = False
    objective.mgSmoothingIterations = 2
    oco = pyVoxelFEM.OCOptimizer(top)
    top.setVars(tps.getDensities())
    history = []
    for idx in range(max_iter):
        history.append(2.0 * top.evaluateObjective())
        mfw = mesh.MSHFieldWriter(str(tmp_path / ("it%d_gt.mesh" % idx)), *tps.getMesh())   # utils.py:324-325
        mfw.addField("density", tps.getDensities())
        oco.step()
    del mfw
    assert history[0] > history[1] > history[2] > 0
    assert abs(top.evaluateConstraints()[0]) <= 1e-6
    x0 = tps.getDensities()
    # utils.compute_binary_compliance_loss (fem.py:98): setVars(binary) / evaluateObjective / setVars(density)
    top.setVars(((x0 > 0.5) * 1.0).astype(np.float64))
    assert np.isfinite(2.0 * top.evaluateObjective())
    top.setVars(x0)
    # checkpoint read-back: utils.load_mesh (MSHFieldParser3.scalarField) and TPS.readDensities
    path = str(tmp_path / "final.msh")
    mesh.MSHFieldWriter(path, *tps.getMesh()).addField("density", x0)
    np.testing.assert_array_equal(mesh.MSHFieldParser3(mshPath=path).scalarField("density"), x0)
    tps2 = pyVoxelFEM.TensorProductSimulator(orderFEM, domainCorners, gridDimensions)
    tps2.readDensities(path)
    np.testing.assert_array_equal(tps2.getDensities(), x0)
    V, F = tps.getMesh()
    assert V.shape == (tps.numNodes(), 3) and F.shape == (tps.numElements(), 8)
    np.testing.assert_array_equal(V[5], tps.nodePosition(5))
    e = tps.numElements() // 2
    ev = [tps.elemNodeGlobalIndex(e, n) for n in range(8)]
    np.testing.assert_array_equal(F[e], [ev[0], ev[1], ev[3], ev[2], ev[4], ev[5], ev[7], ev[6]])   # TPS.hh:547-561
    if helpers is not None:
        assert nonLinearProblem.n == tps.numElements()


@pytest.mark.skipif(not os.path.isdir(REF_HELPERS), reason="the reference tree is only mounted in the build container")
def test_reference_helper_modules_import_against_the_shims(shim_path):
    """helpers/ipopt_helpers.py + mangled_names.py + history_helpers.py of the reference import cleanly on top of the shims
    and resolve the pyVoxelFEM names they touch at import time (no device needed)."""
    import pyVoxelFEM
    helpers = _reference_helpers()
    import mangled_names
    assert mangled_names.FiltersNames == (pyVoxelFEM.ProjectionFilter, pyVoxelFEM.SmoothingFilter, pyVoxelFEM.LangelaarFilter)
    assert mangled_names.ConstraintsNames == (pyVoxelFEM.TotalVolumeConstraint,)
    assert callable(helpers.initializeTensorProductSimulator) and callable(helpers.initializeIpoptProblem)
    with pytest.raises(RuntimeError, match="no CUDA device|CUDA"):
        if pyVoxelFEM.TensorProductSimulator([1, 1], [[0, 0], [1, 1]], [4, 4]):
            raise RuntimeError("CUDA device present")  # on a GPU box the constructor succeeds: still fine


@pytest.mark.gpu
def test_benchmark_sections_carry_the_reference_names(capsys):
    """after benchmark_reset() the bound calls record the reference's own section names (TPS.hh:907, MG.hh:693, Problem.hh:43-132,
    OptimalityCriterion.hh:42); before it nothing is recorded (no device synchronisation in an unprofiled run)"""
    import numpy as np
    import ndr_b200.pyVoxelFEM as V
    from ndr_b200 import benchmark as B
    B._enabled = False; B._sections.clear()
    tps = V.TensorProductSimulator([1, 1, 1], [np.zeros(3), np.array([2.0, 1.0, 1.0])], [16, 8, 8])
    u = np.random.default_rng(0).standard_normal((tps.numNodes(), 3))
    tps.applyK(u)
    assert B.to_dict() == {}
    V.benchmark_reset()
    tps.applyK(u)
    tps.applyK(u)
    d = B.to_dict()
    assert d["applyK 16x8x8"]["calls"] == 2 and d["applyK 16x8x8"]["time"] > 0
    V.benchmark_report()
    assert "applyK 16x8x8" in capsys.readouterr().out
    B._enabled = False; B._sections.clear()
