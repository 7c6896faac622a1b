"""GPU parity: every stage of the device solve (through the C ABI) against the CPU oracle on the same
seeded inputs.  FP64 everywhere; tolerances are relative to the field's max-abs and written per test."""
import os

import numpy as np
import pytest

from conftest import MATERIAL, bc_path

pytestmark = pytest.mark.gpu

CASES = {
    # name: (dim, domain max corner, grid, bc file, levels)
    "cant3d": (3, [2, 1, 1], [32, 16, 16], "3d/cantilever_flexion.bc", 3),
    "bridge3d": (3, [4, 2, 1], [32, 16, 16], "3d/bridge.bc", 2),   # anisotropic element (1/8, 1/8, 1/16)
    "mbb2d": (2, [3, 1], [96, 32], "2d/mbb_beam.bc", 3),
}
# coarsest level 12x12x12: planes of 13*13*3 = 507 dofs (> 256) take the grid-wide chain kernel of coarse_solve.cuh
BIG_COARSE = (3, [1, 1, 1], [24, 24, 24], "3d/cantilever_flexion.bc", 1)


def rel(a, b):
    return float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(b)), 1e-300))


def make_pair(O, name, rho_mode="random"):
    import ndr_b200.pyVoxelFEM as V
    dim, dom, grid, bc, levels = CASES[name]
    o = O.OracleSim([0] * dim, dom, grid)
    o.read_material(MATERIAL)
    g = V.TensorProductSimulator([1] * dim, [np.zeros(dim), np.array(dom, float)], grid)
    g.readMaterial(MATERIAL)
    ne = o.ne
    rho = np.random.default_rng(88).uniform(0, 1, ne) if rho_mode == "random" else np.full(ne, 0.4)
    o.set_densities(rho)
    g.setElementDensities(rho)
    o.apply_bc_file(bc_path(bc))
    g.applyDisplacementsAndLoadsFromFile(bc_path(bc))
    o.set_simp(1.0, 1e-4, 3.0)
    g.E_0, g.E_min, g.gamma = 1.0, 1e-4, 3.0
    return o, g, levels


@pytest.mark.parametrize("name", list(CASES))
def test_setup_parity(oracle_mod, name):
    """K0, node numbering, Dirichlet masks and load vector are identical (indexing + BC handling)."""
    o, g, _ = make_pair(oracle_mod, name)
    assert g.numNodes() == o.nn and g.numElements() == o.ne
    assert rel(g.fullDensityElementStiffnessMatrix(), o.K0()) < 1e-14
    np.testing.assert_array_equal(g.buildLoadVector(), o.load_vector())
    o.mg_create(1)
    np.testing.assert_array_equal(g.dirichletMask, o.mg_level_dmask(0).astype(bool))
    for ni in (0, o.nn // 3, o.nn - 1):
        np.testing.assert_array_equal(g.nodePosition(ni), o.node_position(ni))
    np.testing.assert_array_equal(g.getDensities(), o.densities())


@pytest.mark.parametrize("name", list(CASES))
def test_apply_k_and_gradient(oracle_mod, name):
    o, g, _ = make_pair(oracle_mod, name)
    u = np.random.default_rng(89).standard_normal((o.nn, o.dim))
    assert rel(g.applyK(u), o.apply_k(u)) < 1e-13
    assert rel(g.complianceGradient(u), o.compliance_gradient(u)) < 1e-13


@pytest.mark.parametrize("name", list(CASES))
def test_hierarchy_operators(oracle_mod, name):
    """Dirichlet coarsening, Galerkin operators, smoother, transfers and residual on every level."""
    o, g, levels = make_pair(oracle_mod, name)
    o.mg_create(levels)
    mg = g.multigridSolver(levels)
    o.mg_update_stiffness(); o.mg_update_blockks()
    mg.updateElementStiffnessMatrices()
    rng = np.random.default_rng(7)
    for l in range(levels + 1):
        nn = o.mg_level_nodes(l)
        assert mg.getSimulator(l).numNodes() == nn
        np.testing.assert_array_equal(mg.getSimulator(l).dirichletMask, o.mg_level_dmask(l).astype(bool))
        u, b = rng.standard_normal((nn, o.dim)), rng.standard_normal((nn, o.dim))
        assert rel(mg.applyK(l, u), o.mg_apply_k(l, u)) < 1e-13, l
        assert rel(mg.computeResidual(l, u, b), o.mg_residual(l, u, b)) < 1e-13, l
        np.testing.assert_array_equal(mg.zeroOutDirichletComponents(l, u), o.mg_zero_dirichlet(l, u))
        if 1 <= l < levels:  # oracle assembles block operators for levels 1..L-1
            S = np.moveaxis(mg.levelStencil(l), 3, 0)  # (nn, 3^N, N, N)
            assert rel(S, o.mg_level_stencil(l)) < 1e-13, l
        if l < levels:
            for fwd in (True, False):
                assert rel(mg.smoothingMulticoloredGS(l, u, b, fwd), o.mg_smooth(l, u, b, forward=fwd)) < 1e-12, (l, fwd)
            xc = rng.standard_normal((o.mg_level_nodes(l + 1), o.dim))
            assert rel(mg.restriction(l, u), o.mg_restrict(l, u)) < 1e-14
            assert rel(mg.interpolation(l, xc), o.mg_interpolate(l, xc)) < 1e-14
            assert rel(mg.interpolation(l, xc, accumulate_into=u), o.mg_interpolate(l, xc, accum_into=u)) < 1e-14


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("fmg", [False, True])
def test_cycles(oracle_mod, name, fmg):
    """solve(): V-cycles / full multigrid incl. the exact coarsest-level solve."""
    o, g, levels = make_pair(oracle_mod, name)
    o.mg_create(levels)
    mg = g.multigridSolver(levels)
    f = o.load_vector()
    u0 = np.zeros_like(f)
    got = mg.solve(u0, f, 2, 2, fullMultigrid=fmg)
    ref = o.mg_solve(u0, f, 2, 2, fmg=fmg)
    assert rel(got, ref) < 1e-10


@pytest.mark.parametrize("name", list(CASES))
def test_pcg_parity(oracle_mod, name):
    """MG-PCG: identical iterates at fixed iteration counts, identical iteration count at the production
    tolerance, and <= 1e-10 relative displacement / compliance parity for the converged solve."""
    o, g, levels = make_pair(oracle_mod, name)
    o.mg_create(levels)
    mg = g.multigridSolver(levels)
    f = o.load_vector()
    u0 = np.zeros_like(f)
    for iters in (1, 5):
        got = mg.preconditionedConjugateGradient(u0, f, iters, 1e-30, mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True)
        ref, k, _ = o.mg_pcg(u0, f, iters, 1e-30, 1, 2, True)
        assert k == iters and rel(got, ref) < 1e-9, iters
    got, kg, rn = mg.preconditionedConjugateGradient(u0, f, 100, 1e-4, mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True,
                                                     return_info=True)
    ref, ko, hist = o.mg_pcg(u0, f, 100, 1e-4, 1, 2, True)
    assert kg == ko and abs(rn - hist[-1]) <= 1e-6 * hist[-1]
    assert rel(got, ref) < 1e-9
    got, kg, _ = mg.preconditionedConjugateGradient(u0, f, 100, 1e-12, mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True,
                                                    return_info=True)
    ref, ko, _ = o.mg_pcg(u0, f, 100, 1e-12, 1, 2, True)
    assert rel(got, ref) < 1e-10
    assert abs(np.sum(f * got) - np.sum(f * ref)) <= 1e-10 * abs(np.sum(f * ref))


def test_large_coarsest_planes(oracle_mod):
    """exact coarsest solve with planes above the cluster kernel's size (cooperative grid kernel) inside FMG / PCG"""
    CASES["bigcoarse"] = BIG_COARSE
    try:
        o, g, levels = make_pair(oracle_mod, "bigcoarse")
    finally:
        del CASES["bigcoarse"]
    o.mg_create(levels)
    mg = g.multigridSolver(levels)
    f = o.load_vector()
    u0 = np.zeros_like(f)
    assert rel(mg.solve(u0, f, 2, 2, fullMultigrid=True), o.mg_solve(u0, f, 2, 2, fmg=True)) < 1e-10
    got, kg, _ = mg.preconditionedConjugateGradient(u0, f, 100, 1e-10, mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True, return_info=True)
    ref, ko, _ = o.mg_pcg(u0, f, 100, 1e-10, 1, 2, True)
    assert kg == ko and rel(got, ref) < 1e-9
