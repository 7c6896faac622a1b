#!/usr/bin/env python
"""BASELINE configs[4]: the neural design loop - a Fourier-feature MLP predicts the densities of a 256x128x128 grid, VoxelFEM
returns compliance and per-voxel gradient, Adam updates the network - one optimisation step timed end to end on ONE GPU.

The network is stock PyTorch and NOT part of the product (north_star: "PyTorch appears only in the existing Python
neural-field path"); this file holds a fresh nn.Module of the architecture the reference trains (networks.py:128-185:
gamma(x) = [sin(2 pi B x), cos(2 pi B x)], B ~ N(0, scale^2), Linear/ReLU stack, Sigmoid output; training/train_xdg.py:186-191,
282-331 for the step) so that the loop can run on a box without /root/reference.  The reference keeps the solver on the CPU and
moves densities through numpy (train_xdg.py:302-305); here the densities never leave the GPU (ndr_b200/fem_torch.py).
Usage: python tools/neural_loop.py [EX EY EZ] [steps]"""
import math
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
FIX = os.path.join(ROOT, "tests", "golden", "fixtures")


def build_model(torch, embedding_size=256, n_neurons=256, n_layers=4, scale=2.0, seed=88):
    nn = torch.nn

    class FourierFeatureMLP(nn.Module):
        def __init__(self):
            super().__init__()
            g = torch.Generator().manual_seed(seed)
            self.register_buffer("B", torch.randn(embedding_size, 3, generator=g) * scale)
            layers, width = [], 2 * embedding_size
            for i in range(n_layers):
                last = i == n_layers - 1
                layers.append(nn.Linear(width, 1 if last else n_neurons))
                layers.append(nn.Sigmoid() if last else nn.ReLU())
                width = n_neurons
            self.net = nn.Sequential(*layers)

        def forward(self, coords):
            p = (2.0 * math.pi * coords) @ self.B.T
            return self.net(torch.cat([torch.sin(p), torch.cos(p)], dim=-1))

    return FourierFeatureMLP()


def run(grid, steps=3, levels=3, vol=0.5, chunk=1 << 20, verbose=False):
    import torch
    import ndr_b200.pyVoxelFEM as V
    from ndr_b200.fem_torch import VoxelFEMFunction
    dev = torch.device("cuda")
    tps = V.TensorProductSimulator([1, 1, 1], [np.zeros(3), np.array([2.0, 1.0, 1.0])], grid)
    tps.readMaterial(os.path.join(FIX, "materials", "B9Creator.material"))
    tps.setUniformDensities(vol)
    tps.applyDisplacementsAndLoadsFromFile(os.path.join(FIX, "bcs", "3d", "cantilever_flexion.bc"))
    tps.E_0, tps.E_min, tps.gamma = 1.0, 1e-4, 3.0
    objective = V.MultigridComplianceObjective(tps.multigridSolver(levels))
    objective.tol, objective.mgIterations, objective.fullMultigrid, objective.zeroInit, objective.mgSmoothingIterations = 1e-4, 1, True, False, 2
    top = V.TopologyOptimizationProblem(tps, objective, [V.TotalVolumeConstraint(vol)], [])   # train_xdg.py:267: no VoxelFEM-side filters
    model = build_model(torch).to(dev)
    optim = torch.optim.Adam(model.parameters(), lr=3e-4)
    axes = [(torch.arange(n, device=dev, dtype=torch.float32) + 0.5) / n for n in grid]
    coords = torch.stack(torch.meshgrid(*axes, indexing="ij"), dim=-1).reshape(-1, 3)     # element centres, NDVector order
    ne = coords.shape[0]

    def step():
        optim.zero_grad(set_to_none=True)
        # forward in chunks with recomputation in backward would be the memory-lean form; the B200 holds the whole graph
        density = torch.clamp(model(coords).view(-1), 0.0, 1.0)
        compliance = VoxelFEMFunction.apply(density, top)
        volume_loss = 1500.0 * torch.clamp(density.mean() - vol, min=0.0) ** 2
        loss = compliance + volume_loss
        loss.backward()
        optim.step()
        return float(loss.detach()), float(compliance.detach())

    step()  # warm-up: allocator, graph capture of the preconditioner, cuBLAS handles of torch
    torch.cuda.synchronize()
    times, hist = [], []
    for _ in range(steps):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        l, c = step()
        torch.cuda.synchronize(); times.append((time.perf_counter() - t0) * 1e3)
        hist.append(c)
        its, rn = objective.lastSolve()
        if verbose:
            print("step: %.1f ms, loss %.4f, compliance %.4f, PCG iterations %d" % (times[-1], l, c, its), flush=True)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    x64 = torch.clamp(model(coords).view(-1), 0, 1).detach().double()
    torch.cuda.synchronize(); ev0.record()
    top.setVars(x64, forceUpdate=True); J = top.evaluateObjective(); g = torch.empty_like(x64); top.evaluateObjectiveGradient(out=g)
    ev1.record(); torch.cuda.synchronize()
    return {"grid": grid, "elements": int(ne), "network": "Fourier-feature MLP 3 -> 512 -> 256 x 3 -> 1 (networks.py:128-185), fp32, Adam",
            "ms_per_step": float(np.median(times)), "ms_per_step_all": times, "ms_voxelfem_part": ev0.elapsed_time(ev1),
            "compliance_history": hist, "pcg_iterations_last": int(its),
            "hand_off": "densities and gradient stay on the GPU (CUDA tensors used in place through __cuda_array_interface__)"}


if __name__ == "__main__":
    grid = [int(a) for a in sys.argv[1:4]] if len(sys.argv) >= 4 else [256, 128, 128]
    steps = int(sys.argv[4]) if len(sys.argv) >= 5 else 3
    import json
    print(json.dumps(run(grid, steps, verbose=True)))
