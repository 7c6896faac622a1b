"""Slab-decomposed solve on ONE GPU: W ranks as W host threads sharing the device (ThreadComm, csrc/comm.cu).
Every rank's local results are compared with the CPU oracle's solution of the UNDECOMPOSED grid."""
import threading

import numpy as np
import pytest

from conftest import MATERIAL, bc_path

pytestmark = pytest.mark.gpu

DOM, BC = [2.0, 1.0, 1.0], "3d/cantilever_flexion.bc"


def rel(a, b):
    return float(np.max(np.abs(np.asarray(a) - np.asarray(b))) / max(np.max(np.abs(b)), 1e-300))


def oracle_problem(O, grid, levels):
    o = O.OracleSim([0, 0, 0], DOM, grid)
    o.read_material(MATERIAL)
    rho = np.random.default_rng(88).uniform(0, 1, o.ne)
    o.set_densities(rho)
    o.apply_bc_file(bc_path(BC))
    o.set_simp(1.0, 1e-4, 3.0)
    o.mg_create(levels)
    return o, rho


def run_ranks(world, body):
    """runs body(rank) on `world` threads; re-raises the first failure"""
    errs = [None] * world

    def wrap(r):
        try:
            body(r)
        except BaseException as e:  # noqa: BLE001
            errs[r] = e
    ts = [threading.Thread(target=wrap, args=(r,)) for r in range(world)]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    bad = [(r, e) for r, e in enumerate(errs) if e is not None]
    if bad:
        # a rank that fails leaves the others waiting at a collective until they time out: report the root cause first
        bad.sort(key=lambda re: "did not reach the collective" in str(re[1]))
        raise RuntimeError("; ".join("rank %d: %s: %s" % (r, type(e).__name__, e) for r, e in bad)) from bad[0][1]


def make_rank(grid, levels, rank, world, comms, hs, rho):
    from ndr_b200.slab import SlabSolver
    s = SlabSolver(grid, np.zeros(3), np.array(DOM), levels, rank, world, comms[rank], hs.bind(rank), MATERIAL)
    s.tps.setElementDensities(np.ascontiguousarray(rho[s.layout.elem_slice()]))
    s.apply_bc_file(bc_path(BC))
    s.tps.E_0, s.tps.E_min, s.tps.gamma = 1.0, 1e-4, 3.0
    return s


@pytest.mark.parametrize("grid,levels,world", [([32, 16, 16], 2, 2), ([32, 16, 16], 2, 4), ([48, 16, 16], 3, 3), ([32, 8, 8], 1, 5), ([64, 32, 32], 3, 2)])
def test_slab_setup_and_apply(oracle_mod, grid, levels, world):
    """local load vector / Dirichlet mask = windows of the global ones; K u on the owned planes = global K u"""
    from ndr_b200.slab import ThreadHostSum, make_thread_comms
    o, rho = oracle_problem(oracle_mod, grid, levels)
    f, dm = o.load_vector(), o.mg_level_dmask(0).astype(bool)
    u = np.random.default_rng(89).standard_normal((o.nn, 3))
    Ku = o.apply_k(u)
    comms, hs = make_thread_comms(world), ThreadHostSum(world)

    def body(r):
        s = make_rank(grid, levels, r, world, comms, hs, rho)
        L = s.layout
        np.testing.assert_array_equal(s.tps.buildLoadVector(), f[L.node_slice()])
        np.testing.assert_array_equal(s.tps.dirichletMask, dm[L.node_slice()])
        out = s.tps.applyK(np.ascontiguousarray(u[L.node_slice()]))
        assert rel(out[L.owned_node_rows()], Ku[L.owned_global_node_rows()]) < 1e-13
    run_ranks(world, body)


@pytest.mark.parametrize("grid,levels,world", [([32, 16, 16], 2, 2), ([32, 16, 16], 2, 4), ([48, 16, 16], 3, 3)])
def test_slab_hierarchy_and_smoother(oracle_mod, grid, levels, world):
    """per-level Dirichlet masks, one coloured GS sweep and the residual on every level against the oracle's"""
    from ndr_b200.slab import ThreadHostSum, make_thread_comms
    o, rho = oracle_problem(oracle_mod, grid, levels)
    o.mg_update_stiffness()
    rng = np.random.default_rng(5)
    ref = {}
    for l in range(levels):
        nn = o.mg_level_nodes(l)
        u, b = rng.standard_normal((nn, 3)), rng.standard_normal((nn, 3))
        u = o.mg_zero_dirichlet(l, u)
        ref[l] = (u, b, o.mg_smooth(l, u.copy(), b, True), o.mg_smooth(l, u.copy(), b, False), o.mg_residual(l, u, b), o.mg_level_dmask(l).astype(bool),
                  o.apply_k(u) if l == 0 else None)
    comms, hs = make_thread_comms(world), ThreadHostSum(world)

    def body(r):
        import ctypes as C
        from ndr_b200._lib import check, lib
        s = make_rank(grid, levels, r, world, comms, hs, rho)
        mg = s.multigrid()
        pn = s.layout.plane_nodes
        for l in range(levels):
            u, b, uf, ub, res, dm, Ku = ref[l]
            w = 1 << l
            first, last = C.c_size_t(), C.c_size_t()
            check(lib().vfem_mg_owned_planes(mg._h, l, C.byref(first), C.byref(last)))
            pnl = (grid[1] // w + 1) * (grid[2] // w + 1)
            p0 = s.layout.layer_lo // w
            nloc = mg.getSimulator(l).numNodes()
            win = slice(p0 * pnl, p0 * pnl + nloc)
            own_l = slice(first.value * pnl, (last.value + 1) * pnl)
            own_g = slice((p0 + first.value) * pnl, (p0 + last.value + 1) * pnl)
            if l > 0:
                np.testing.assert_array_equal(mg.getSimulator(l).dirichletMask[own_l], dm[own_g])
            for fwd, want in ((True, uf), (False, ub)):
                got = mg.smoothingMulticoloredGS(l, np.ascontiguousarray(u[win]), np.ascontiguousarray(b[win]), fwd)
                assert rel(got[own_l], want[own_g]) < 1e-12, (l, fwd)
            got = mg.computeResidual(l, np.ascontiguousarray(u[win]), np.ascontiguousarray(b[win]))
            assert rel(got[own_l], res[own_g]) < 1e-12, l
            if l == 0:  # ghost refresh + K-apply in one call (halo shift behind the interior layers; three disjoint views)
                import torch
                ul = torch.from_numpy(np.ascontiguousarray(u[win])).cuda()
                ghost = np.ones(nloc, bool); ghost[own_l] = False
                ul[torch.from_numpy(ghost).cuda()] = 7.0          # stale ghost planes: the call must refresh them
                out = torch.full_like(ul, np.nan)
                check(lib().vfem_mg_apply_k_halo(mg._h, C.c_void_p(ul.data_ptr()), C.c_void_p(out.data_ptr()), 1))
                torch.cuda.synchronize()
                assert rel(out.cpu().numpy()[own_l], Ku[own_g]) < 1e-13
    run_ranks(world, body)


@pytest.mark.parametrize("grid,levels,world", [([32, 16, 16], 2, 2), ([32, 16, 16], 2, 4), ([48, 16, 16], 3, 3), ([64, 16, 16], 3, 8),
                                               ([64, 32, 32], 3, 2), ([128, 16, 16], 2, 4)])   # the last two: 32 element layers per rank
def test_slab_mg_pcg(oracle_mod, grid, levels, world):
    """MG-PCG (FMG preconditioner, 2 sweeps): fixed-iteration iterates and the converged solution against the oracle"""
    from ndr_b200.slab import ThreadHostSum, make_thread_comms
    o, rho = oracle_problem(oracle_mod, grid, levels)
    f = o.load_vector()
    want = {}
    for its, tol in ((3, 1e-30), (100, 1e-10)):
        want[its] = o.mg_pcg(np.zeros_like(f), f, its, tol, 1, 2, True)
    comms, hs = make_thread_comms(world), ThreadHostSum(world)

    def body(r):
        s = make_rank(grid, levels, r, world, comms, hs, rho)
        L = s.layout
        mg = s.multigrid()
        fl = s.tps.buildLoadVector()
        for its, tol in ((3, 1e-30), (100, 1e-10)):
            uo, no, _ = want[its]
            u, n, _ = mg.preconditionedConjugateGradient(np.zeros_like(fl), fl, its, tol, mgIterations=1, mgSmoothingIterations=2,
                                                         fullMultigrid=True, return_info=True)
            assert n == no
            scale = np.max(np.abs(uo))
            # owned planes AND the ghost planes next to them (the returned field is consistent)
            lo, hi = max(L.own0 - 1, 0), min(L.own1 + 1, L.local_planes - 1)
            rows = slice(lo * L.plane_nodes, (hi + 1) * L.plane_nodes)
            grow = slice((L.layer_lo + lo) * L.plane_nodes, (L.layer_lo + hi + 1) * L.plane_nodes)
            assert np.max(np.abs(u[rows] - uo[grow])) / scale < 1e-10, (its, r)
        # compliance 1/2 f.u over the owned dofs of all ranks
        J = 0.5 * s.dot(fl, u)
        assert abs(J - 0.5 * float(np.sum(f * want[100][0]))) < 1e-10 * abs(J)
    run_ranks(world, body)


@pytest.mark.parametrize("grid,levels,world", [([32, 16, 16], 2, 2), ([32, 16, 16], 2, 4), ([48, 16, 16], 3, 3)])
def test_slab_topology_optimization_iteration(oracle_mod, grid, levels, world):
    """The density pipeline on slabs (SURVEY 8e "rho/E ghost element layer"): box filter with a layer halo + global clipping,
    projection, volume constraint (owned sums + all-reduce), objective / gradient chain and three optimality-criterion steps
    against the oracle's undecomposed problem: lambda equal, max-abs density <= 1e-8 on the owned layers of every rank."""
    import ndr_b200.pyVoxelFEM as V
    from ndr_b200.slab import ThreadHostSum, make_thread_comms
    vol = 0.4
    o, _ = oracle_problem(oracle_mod, grid, levels)
    x0g = np.random.default_rng(11).uniform(0.25, 0.6, o.ne)
    o.set_densities(x0g)
    o.top_create(True, vol, [("smoothing", 1), ("projection", 1.0)])
    o.top_set_solver(100, 1e-11, 1, 2, True, False)
    o.top_set_vars(x0g)
    ref = {"dens": o.top_vars(-1).copy(), "J": o.top_objective(), "grad": o.top_gradient().copy(), "c": o.top_constraint(),
           "jac": o.top_constraint_jac().copy(), "steps": []}
    for _ in range(3):
        lam, _ = o.oc_step()
        ref["steps"].append((lam, o.top_vars(0).copy(), o.top_objective()))
    comms, hs = make_thread_comms(world), ThreadHostSum(world)

    def body(r):
        s = make_rank(grid, levels, r, world, comms, hs, x0g)
        L = s.layout
        mg = s.multigrid()
        obj = V.MultigridComplianceObjective(mg)
        top = V.TopologyOptimizationProblem(s.tps, obj, [V.TotalVolumeConstraint(vol)], [V.SmoothingFilter(), V.ProjectionFilter()])
        obj.tol = 1e-11
        loc, own, gown = L.elem_slice(), L.owned_elem_rows(), L.owned_global_elem_rows()
        xl = x0g[loc].copy()                          # a private copy: the ranks are threads sharing x0g
        xl[: own.start] = 0.123                       # stale ghost layers: setVars must refresh them from the owner
        assert top.setVars(xl)
        assert rel(top.getDensities()[own], ref["dens"][gown]) < 1e-14
        np.testing.assert_allclose(top.getDensities(), ref["dens"][loc], rtol=0, atol=1e-14)   # ghost layers consistent too
        assert abs(top.evaluateObjective() - ref["J"]) <= 1e-9 * abs(ref["J"])
        assert rel(top.evaluateObjectiveGradient()[own], ref["grad"][gown]) < 1e-8
        assert abs(top.evaluateConstraints()[0] - ref["c"]) < 1e-13
        assert rel(top.evaluateConstraintsJacobian()[0][own], ref["jac"][gown]) < 1e-13
        oc = V.OCOptimizer(top)
        for k, (lam, xg, J) in enumerate(ref["steps"]):
            oc.step()
            assert abs(oc.lastLambda - lam) <= 1e-9 * lam, k
            assert np.max(np.abs(top.getVars()[own] - xg[gown])) <= 1e-8, k
            assert abs(top.evaluateObjective() - J) <= 1e-8 * J, k
    run_ranks(world, body)


@pytest.mark.parametrize("grid,levels,world", [([64, 32, 16], 2, 2), ([64, 32, 16], 3, 4), ([64, 32, 16], 3, 8)])
def test_slab_mg_pcg_bridge(oracle_mod, grid, levels, world):
    """BASELINE config 3's boundary conditions on slabs (bridge.bc: a clamped face on rank 0, an x-roller face on the LAST
    rank, a force region that spans every rank): same iteration count as the oracle's undecomposed solve at the production
    tolerance, converged solution equal"""
    from ndr_b200.slab import SlabSolver, ThreadHostSum, make_thread_comms
    dom, bc = [4.0, 2.0, 1.0], bc_path("3d/bridge.bc")
    o = oracle_mod.OracleSim([0, 0, 0], dom, grid)
    o.read_material(MATERIAL)
    rho = np.random.default_rng(88).uniform(0, 1, o.ne)
    o.set_densities(rho)
    o.apply_bc_file(bc)
    o.set_simp(1.0, 1e-4, 3.0)
    o.mg_create(levels)
    f, dm = o.load_vector(), o.mg_level_dmask(0).astype(bool)
    want = {tol: o.mg_pcg(np.zeros_like(f), f, 100, tol, 1, 2, True) for tol in (1e-4, 1e-10)}
    comms, hs = make_thread_comms(world), ThreadHostSum(world)

    def body(r):
        s = SlabSolver(grid, np.zeros(3), np.array(dom), levels, r, world, comms[r], hs.bind(r), MATERIAL)
        L = s.layout
        s.tps.setElementDensities(rho[L.elem_slice()].copy())
        s.apply_bc_file(bc)
        s.tps.E_0, s.tps.E_min, s.tps.gamma = 1.0, 1e-4, 3.0
        np.testing.assert_array_equal(s.tps.buildLoadVector(), f[L.node_slice()])
        np.testing.assert_array_equal(s.tps.dirichletMask, dm[L.node_slice()])
        mg = s.multigrid()
        fl = s.tps.buildLoadVector()
        for tol in (1e-4, 1e-10):
            uo, no, _ = want[tol]
            u, n, _ = mg.preconditionedConjugateGradient(np.zeros_like(fl), fl, 100, tol, mgIterations=1, mgSmoothingIterations=2,
                                                         fullMultigrid=True, return_info=True)
            assert n == no, (tol, n, no)
            if tol == 1e-10:
                rows, grow = L.owned_node_rows(), L.owned_global_node_rows()
                assert np.max(np.abs(u[rows] - uo[grow])) / np.max(np.abs(uo)) < 1e-8
    run_ranks(world, body)
