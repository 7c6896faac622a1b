"""GPU parity of the element-marching Walsh-Hadamard K-apply (ndr_b200/csrc/apply_march.cuh) against the CPU oracle
and against the library's own per-node gather kernel, over grid shapes that exercise every tile / ghost / special
row-column path (tile shapes 8x64, 16x32, 15x33, 7x65; one and several chunks along the march)."""
import numpy as np
import pytest

from conftest import MATERIAL

pytestmark = pytest.mark.gpu

GRIDS = [
    ([5, 4, 6], [1, 1, 1]), ([3, 9, 31], [1, 2, 1]), ([4, 8, 32], [2, 1, 1]), ([6, 17, 33], [1, 1, 1]),
    ([9, 7, 63], [1, 1, 2]), ([7, 16, 64], [4, 2, 1]), ([5, 15, 65], [1, 1, 1]), ([3, 33, 97], [1, 1, 1]),
    ([40, 20, 10], [4, 2, 1]), ([2, 2, 2], [1, 1, 1]), ([1, 1, 1], [1, 1, 1]), ([1, 40, 70], [1, 1, 1]),
    ([32, 16, 16], [2, 1, 1]), ([24, 64, 64], [2, 1, 1]), ([13, 30, 128], [1, 1, 1]), ([10, 66, 129], [3, 1, 2]),
]


def rel(a, b):
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-300))


def make(grid, dom):
    import ndr_b200.pyVoxelFEM as V
    g = V.TensorProductSimulator([1, 1, 1], [np.zeros(3), np.array(dom, float)], grid)
    g.readMaterial(MATERIAL)
    rho = np.random.default_rng(88).uniform(0, 1, int(np.prod(grid)))
    g.setElementDensities(rho)
    g.E_0, g.E_min, g.gamma = 1.0, 1e-4, 3.0
    return g, rho


@pytest.mark.parametrize("grid,dom", GRIDS)
def test_march_matches_gather_and_oracle(oracle_mod, grid, dom):
    g, rho = make(grid, dom)
    info = g.applyInfo()
    assert info["march"] == 1 and info["k0_structure"] == 2, info
    u = np.random.default_rng(89).standard_normal((g.numNodes(), 3))
    a = g.applyK(u)
    assert np.isfinite(a).all()
    g.setApplyVariant(1)
    assert g.applyInfo()["march"] == 0
    b = g.applyK(u)
    g.setApplyVariant(0)
    assert rel(a, b) < 1e-13, (info, rel(a, b))
    if np.prod(grid) <= 70000:
        o = oracle_mod.OracleSim([0, 0, 0], dom, grid)
        o.read_material(MATERIAL); o.set_densities(rho); o.set_simp(1.0, 1e-4, 3.0)
        assert rel(a, o.apply_k(u)) < 1e-13
    # bit-reproducible: same launch, same answer
    np.testing.assert_array_equal(a, g.applyK(u))


@pytest.mark.parametrize("shape", range(13))
@pytest.mark.parametrize("xc", [None, 4])
def test_every_tile_shape(shape, xc, monkeypatch):
    """each compiled tile shape forced through the planner overrides (several tiles along y and z, several chunks), all
    three store modes, against the per-node gather kernel"""
    monkeypatch.setenv("VFEM_MARCH_SHAPE", str(shape))
    if xc:
        monkeypatch.setenv("VFEM_MARCH_XC", str(xc))
    from conftest import bc_path
    grid, dom = [10, 24, 72], [1, 2, 6]
    g, _ = make(grid, dom)
    g.applyDisplacementsAndLoadsFromFile(bc_path("3d/cantilever_flexion.bc"))
    info = g.applyInfo()
    assert info["march"] == 1 and info["tile_shape"] == shape and (xc is None or info["layers_per_chunk"] == xc), info
    rng = np.random.default_rng(shape)
    u, b = rng.standard_normal((g.numNodes(), 3)), rng.standard_normal((g.numNodes(), 3))
    mg = g.multigridSolver(1)
    a, r = g.applyK(u), mg.computeResidual(0, u, b)
    g.setApplyVariant(1)
    assert g.applyInfo()["march"] == 0
    assert rel(a, g.applyK(u)) < 1e-13 and rel(r, mg.computeResidual(0, u, b)) < 1e-13


def test_march_residual_and_masked_modes(oracle_mod):
    """MODE 1 (mask(b - K u)) and MODE 2 (mask(K u)) through MultigridSolver.computeResidual / PCG's A d."""
    import ndr_b200.pyVoxelFEM as V
    from conftest import bc_path
    grid, dom = [16, 16, 64], [2, 1, 1]
    g, rho = make(grid, dom)
    g.applyDisplacementsAndLoadsFromFile(bc_path("3d/cantilever_flexion.bc"))
    mg = g.multigridSolver(2)
    rng = np.random.default_rng(5)
    u, b = rng.standard_normal((g.numNodes(), 3)), rng.standard_normal((g.numNodes(), 3))
    r1 = mg.computeResidual(0, u, b)
    g.setApplyVariant(1)
    r2 = mg.computeResidual(0, u, b)
    g.setApplyVariant(0)
    assert rel(r1, r2) < 1e-13
    mask = g.dirichletMask
    assert mask.any() and np.all(r1[mask] == 0.0)


def test_dense_block_variant_matches():
    """The variant for a K0 without the isotropic zero pattern (dense 3x3 blocks) gives the same answer."""
    g, _ = make([6, 9, 40], [1, 2, 1])
    u = np.random.default_rng(1).standard_normal((g.numNodes(), 3))
    g.setApplyVariant(2)
    a = g.applyK(u)
    g.setApplyVariant(1)
    assert rel(a, g.applyK(u)) < 1e-13


def test_large_grid_linearity_and_symmetry():
    """Full BASELINE size (128x64x64): size-independent properties — linearity and <v, K u> = <u, K v>."""
    g, _ = make([128, 64, 64], [2, 1, 1])
    rng = np.random.default_rng(7)
    u, v = rng.standard_normal((g.numNodes(), 3)), rng.standard_normal((g.numNodes(), 3))
    Ku, Kv = g.applyK(u), g.applyK(v)
    assert rel(g.applyK(2.0 * u - 3.0 * v), 2.0 * Ku - 3.0 * Kv) < 1e-13
    assert abs(np.vdot(v, Ku) - np.vdot(u, Kv)) < 1e-11 * abs(np.vdot(v, Ku))
    # rigid translations are in the null space
    t = np.tile(np.array([1.0, -2.0, 0.5]), (g.numNodes(), 1))
    assert np.max(np.abs(g.applyK(t))) < 1e-12 * np.max(np.abs(Ku))
    g.setApplyVariant(1)
    assert rel(Ku, g.applyK(u)) < 1e-13
