"""Neural-design hand-off (fem.py:109-134): torch CUDA densities -> device solve -> per-voxel gradient,
zero-copy, with the reference's loss / gradient conventions; checked against the oracle's objective and gradient."""
import numpy as np
import pytest

from test_gpu_topopt import make_problem
from test_gpu_parity import rel

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["cant3d", "mbb2d"])
def test_voxelfem_function_cuda_and_cpu(oracle_mod, name):
    import torch
    from ndr_b200.fem_torch import VoxelFEMFunction
    o, g, top, obj = make_problem(oracle_mod, name, filters=False)     # config 5: TopologyOptimizationProblem(..., filters=[])
    x = np.random.default_rng(7).uniform(0.2, 0.9, o.ne).astype(np.float32)
    o.top_set_vars(x.astype(np.float64))
    Jo, go = o.top_objective(), o.top_gradient()
    # a tiny "network": densities = sigmoid(w) on the GPU, loss = VoxelFEM compliance
    w = torch.logit(torch.from_numpy(x)).cuda().requires_grad_(True)
    dens = torch.sigmoid(w)
    loss = VoxelFEMFunction.apply(dens, top)
    assert loss.is_cuda and loss.dtype == torch.float32
    xs = dens.detach().double().cpu().numpy()
    o.top_set_vars(xs)
    Jo, go = o.top_objective(), o.top_gradient()
    assert abs(loss.item() - 2.0 * Jo) <= 2e-6 * abs(2.0 * Jo)          # float32 scalar
    loss.backward()
    want = go.astype(np.float32) * (xs * (1 - xs)).astype(np.float32)   # dJ/drho (no factor 2) chained through the sigmoid
    assert rel(w.grad.cpu().numpy(), want) < 1e-5
    # the CPU-tensor path is the reference's numpy path
    wc = torch.logit(torch.from_numpy(x)).requires_grad_(True)
    lc = VoxelFEMFunction.apply(torch.sigmoid(wc), top)
    lc.backward()
    assert abs(lc.item() - loss.item()) <= 1e-6 * abs(loss.item())
    assert rel(wc.grad.numpy(), w.grad.cpu().numpy()) < 1e-5


def test_device_arrays_are_used_in_place(oracle_mod):
    """setVars / getDensities / evaluateObjectiveGradient with CUDA tensors: same results as the numpy path"""
    import torch
    o, g, top, obj = make_problem(oracle_mod, "cant3d", filters=True)
    x = np.random.default_rng(9).uniform(0.2, 0.9, o.ne)
    top.setVars(x)
    J, gr, dens = top.evaluateObjective(), top.evaluateObjectiveGradient(), top.getDensities()
    xd = torch.from_numpy(x).cuda()
    assert top.setVars(xd * 0.5 + 0.1)                                  # different variables first
    assert top.setVars(xd)
    assert abs(top.evaluateObjective() - J) <= 1e-9 * abs(J)   # the warm-started PCG (tol 1e-10) restarts from another iterate
    gd = torch.empty_like(xd); dd = torch.empty_like(xd)
    top.evaluateObjectiveGradient(out=gd); top.getDensities(out=dd)
    np.testing.assert_allclose(gd.cpu().numpy(), gr, rtol=0, atol=1e-8 * np.max(np.abs(gr)))
    np.testing.assert_array_equal(dd.cpu().numpy(), dens)


def test_fourier_feature_mlp_design_step():
    """BASELINE configs[4] in small: a Fourier-feature MLP (architecture of networks.py:128-185) drives VoxelFEMFunction on the
    GPU; Adam steps run end to end (tools/neural_loop.py) and the compliance gradient reaches the network weights."""
    import os, sys
    from conftest import ROOT
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import neural_loop
    r = neural_loop.run([32, 16, 16], steps=4, levels=2)
    assert r["elements"] == 32 * 16 * 16 and len(r["compliance_history"]) == 4
    assert all(np.isfinite(r["compliance_history"])) and r["pcg_iterations_last"] > 0
    assert r["compliance_history"][-1] != r["compliance_history"][0]   # the weights moved
