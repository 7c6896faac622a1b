#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native VoxelFEM solve path (BASELINE.json metric:
"K-apply GDOF/s (% HBM roofline); ms per MG-PCG solve").

A step = one matrix-free K(rho) u over the grid (3-D cantilever 128x64x64 per GPU, BASELINE configs[1]).
  value   : GDOF/s of the K-apply with u, K u and E(rho) resident in HBM (device pointers through the C ABI)
  e2e     : the same apply through the reference-facing call `tps.applyK(numpy u) -> numpy` (H2D + D2H inside)
  solve   : ms per MG-PCG solve (reference production settings: FMG preconditioner, 2 smoothing sweeps, tol 1e-4)
  roofline: algorithmic bytes 8*(2*N*nn + ne) per apply / CUDA-event time, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline: the CPU restatement of the reference algorithm (oracle/, OpenMP) on this box's host cores
N > 1 (torchrun): slabs along storage axis 0 (ndr_b200/slab.py); the K-apply step exchanges one node plane per interior
face (ncclSend/ncclRecv issued by the C++ solver on its stream) and `solve` is the slab-decomposed MG-PCG; weak scaling.
`--impl reference` times the CPU restatement (the reference itself cannot be built here: no Eigen/TBB/SuiteSparse).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
FIX = os.path.join(ROOT, "tests", "golden", "fixtures")
MATERIAL = os.path.join(FIX, "materials", "B9Creator.material")
BC = os.path.join(FIX, "bcs", "3d", "cantilever_flexion.bc")
GRID = [128, 64, 64]
DOMAIN = [2.0, 1.0, 1.0]
L2_BYTES = 126e6


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            time.sleep(0.15)
            self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def algorithmic_bytes(nn, ne, N=3):
    return 8 * (2 * N * nn + ne)  # read u, write K u, read E(rho)   


def run_cpu_kapply(budget_s, threads=None, solve=True):
    """CPU restatement of the reference (oracle/): K-apply GDOF/s on the bench grid + one MG-PCG solve."""
    from oracle import oracle as O
    O.build()
    if threads:
        O.set_num_threads(threads)
    s = O.OracleSim([0, 0, 0], DOMAIN, GRID)
    s.read_material(MATERIAL)
    s.set_densities(np.random.default_rng(88).uniform(0, 1, s.ne))
    s.apply_bc_file(BC)
    s.set_simp(1.0, 1e-4, 3.0)
    u = np.random.default_rng(89).standard_normal((s.nn, 3))
    s.apply_k(u)
    t0, reps = time.perf_counter(), 0
    while reps < 3 or (time.perf_counter() - t0 < budget_s and reps < 200):
        s.apply_k(u); reps += 1
    dt = (time.perf_counter() - t0) / reps
    out = {"value": 3 * s.nn / dt / 1e9, "unit": "GDOF/s", "cores": O.max_threads(), "kind": "port",
           "sample": "%d K-applies of the 128x64x64 cantilever (seed-88 densities), OpenMP oracle restatement; the reference binary is "
                     "unbuildable here (Eigen/TBB/SuiteSparse absent)" % reps, "ms_per_apply": dt * 1e3}
    if solve:
        s.mg_create(3)
        f = s.load_vector()
        t0 = time.perf_counter()
        _, its, _ = s.mg_pcg(np.zeros_like(f), f, 100, 1e-4, 1, 2, True)
        out["solve_ms"] = (time.perf_counter() - t0) * 1e3
        out["solve_iterations"] = int(its)
    return out


def reference_arm(args, rank):
    if rank != 0:
        return
    W, K = args.warmup, args.steps
    from oracle import oracle as O
    O.build()
    s = O.OracleSim([0, 0, 0], DOMAIN, GRID)
    s.read_material(MATERIAL)
    s.set_densities(np.random.default_rng(88).uniform(0, 1, s.ne))
    s.set_simp(1.0, 1e-4, 3.0)
    u = np.random.default_rng(89).standard_normal((s.nn, 3))
    for _ in range(W):
        s.apply_k(u)
    t0 = time.perf_counter()
    for _ in range(K):
        s.apply_k(u)
    dt = (time.perf_counter() - t0) / K
    val = 3 * s.nn / dt / 1e9
    cb = {"value": val, "unit": "GDOF/s", "cores": O.max_threads(), "kind": "port",
          "sample": "%d K-applies, 128x64x64 cantilever, OpenMP restatement of the reference algorithm (reference unbuildable here)" % K}
    line = {"impl": "reference", "metric": "k_apply_gdof_per_s", "value": val, "unit": "GDOF/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "3D cantilever 128x64x64 Q1 hex, matrix-free K(rho)u, seed-88 densities"},
            "cpu_baseline": cb, "e2e": {"value": val, "unit": "GDOF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)


def layer_densities(layer_lo, layer_hi, plane_elems):
    """synthetic densities of global element layers [layer_lo, layer_hi): one seeded stream per layer, so that the ghost
    layers of neighbouring ranks hold identical values whatever the decomposition"""
    return np.concatenate([np.random.default_rng([88, k]).uniform(0, 1, plane_elems) for k in range(layer_lo, layer_hi)])


def our_arm(args):
    import torch
    import torch.distributed as dist
    from ndr_b200 import _lib
    from ndr_b200 import slab as S
    import ndr_b200.pyVoxelFEM as V
    from ndr_b200.voxelfem import _dev_ptr

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    _lib.check(_lib.lib().vfem_set_device(local))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, K = args.warmup, args.steps
    hbm_peak, peak_src = measured_peaks()
    L = _lib.lib()

    # ---- workload: weak scaling, one 128x64x64 block of element layers per GPU, slabs along axis 0 ----
    gE = [GRID[0] * world, GRID[1], GRID[2]]
    bbmin, bbmax = np.zeros(3), np.array([DOMAIN[0] * world, DOMAIN[1], DOMAIN[2]])
    if world == 1:
        tps = V.TensorProductSimulator([1, 1, 1], [bbmin, bbmax], gE)
        tps.readMaterial(MATERIAL)
        tps.setElementDensities(np.random.default_rng(88).uniform(0, 1, tps.numElements()))
        tps.applyDisplacementsAndLoadsFromFile(BC)
        tps.E_0, tps.E_min, tps.gamma = 1.0, 1e-4, 3.0
        mg = tps.multigridSolver(args.levels)
        halo_bytes = 0
        apply_halo = None
    else:
        comm = S.make_nccl_comm(dist, rank, world)
        sl = S.SlabSolver(gE, bbmin, bbmax, args.levels, rank, world, comm, S.dist_host_sum(dist, torch.device("cuda", local)), MATERIAL)
        tps, lay = sl.tps, sl.layout
        tps.setElementDensities(layer_densities(lay.layer_lo, lay.layer_hi, lay.plane_elems))
        sl.apply_bc_file(BC)
        tps.E_0, tps.E_min, tps.gamma = 1.0, 1e-4, 3.0
        mg = sl.multigrid()
        halo_bytes = lay.plane_nodes * 24
        apply_halo = lambda u, o: _lib.check(L.vfem_mg_apply_k_halo(mg._h, _dev_ptr(u, u.numel()), _dev_ptr(o, o.numel()), 1))
    nn_l, ne_l = tps.numNodes(), tps.numElements()
    nn_global = (gE[0] + 1) * (gE[1] + 1) * (gE[2] + 1)
    set_bytes = algorithmic_bytes(nn_l, ne_l)
    nsets = max(2, int(np.ceil(2.5 * L2_BYTES / (16 * 3 * nn_l))))  # ring of u/Ku pairs larger than L2
    gen = torch.Generator(device="cuda"); gen.manual_seed(89 + rank)
    us = [torch.randn(nn_l, 3, dtype=torch.float64, device="cuda", generator=gen) for _ in range(nsets)]
    outs = [torch.empty(nn_l, 3, dtype=torch.float64, device="cuda") for _ in range(nsets)]

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(k):  # the local matrix-free apply; N > 1: + the halo shift of u (one node plane per interior face over NVLink),
        # hidden behind the interior element layers
        if apply_halo:
            apply_halo(us[k % nsets], outs[k % nsets])
        else:
            tps.applyK(us[k % nsets], out=outs[k % nsets])

    for k in range(W):
        step(k)
    sync_all()
    sampler = ClockSampler(local); sampler.start()
    launches0 = L.vfem_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync_all()
    ev0.record()
    for k in range(K):
        step(k)
    ev1.record()
    sync_all()
    ms = ev0.elapsed_time(ev1)
    launches = L.vfem_launch_count() - launches0
    # keep the GPU busy long enough for nvidia-smi to see clocks under load (not part of the measurement)
    t_end = time.perf_counter() + 0.6
    while time.perf_counter() < t_end:
        for k in range(50):
            tps.applyK(us[k % nsets], out=outs[k % nsets])
        torch.cuda.synchronize()
    clocks = sampler.stop()
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    value = 3.0 * nn_global * K / (ms * 1e-3) / 1e9

    # ---- dominant kernel alone (no halo exchange): roofline ----
    torch.cuda.synchronize()
    ev0.record()
    for k in range(K):
        tps.applyK(us[k % nsets], out=outs[k % nsets])
    ev1.record(); torch.cuda.synchronize()
    kms = ev0.elapsed_time(ev1) / K
    achieved = set_bytes / (kms * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "applyMarchKernel<0,...> (matrix-free K(rho)u, element-marching Walsh-Hadamard form)", "achieved": achieved,
                "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak, "traffic": None, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": set_bytes, "kernel_ms": kms, "gdof_per_s_per_gpu": 3.0 * nn_l / (kms * 1e-3) / 1e9}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        roofline["traffic"] = json.load(open(tp)).get("apply_march_dram_bytes_per_launch")

    # ---- MG-PCG solve over all ranks (collective: every rank takes part) ----
    solve = None
    if not args.no_solve:
        solve = time_solve(V, torch, args, tps, mg, world, dist)

    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return

    # ---- e2e: the reference-facing call with host buffers (pinned), H2D + D2H inside the timed region ----
    uh = torch.randn(nn_l, 3, dtype=torch.float64).pin_memory().numpy()
    oh = torch.empty(nn_l, 3, dtype=torch.float64).pin_memory().numpy()
    ke = max(3, min(K, 20))
    for _ in range(2):
        tps.applyK(uh, out=oh)
    torch.cuda.synchronize(); ev0.record()
    for _ in range(ke):
        tps.applyK(uh, out=oh)
    ev1.record(); torch.cuda.synchronize()
    e2e_ms = ev0.elapsed_time(ev1) / ke
    e2e = {"value": 3.0 * nn_l * world / (e2e_ms * 1e-3) / 1e9, "unit": "GDOF/s", "h2d_bytes_per_step": int(uh.nbytes), "d2h_bytes_per_step": int(oh.nbytes),
           "ms_per_step": e2e_ms, "call": "pyVoxelFEM TensorProductSimulator.applyK(numpy) -> vfem_tps_apply_k(VFEM_HOST), rank 0's slab"}

    cpu = run_cpu_kapply(args.cpu_seconds, solve=not args.no_solve) if world == 1 and not args.no_cpu else None
    line = {"metric": "k_apply_gdof_per_s", "value": value, "unit": "GDOF/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "3D cantilever %dx%dx%d Q1 hex (128x64x64 element layers per GPU, slabs along axis 0), matrix-free K(rho)u, "
                                   "seed-88 uniform(0,1) densities, E_min=1e-4, p=3" % tuple(gE),
                       "dofs": 3 * nn_global, "l2": "ring of %d u/Ku buffer pairs (%.0f MB) > 126 MB L2" % (nsets, nsets * 48 * nn_l / 1e6),
                       "halo": "none" if world == 1 else "1 node plane (%d B) per interior face per apply, ncclSend/ncclRecv on the solver stream" % halo_bytes},
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline}
    if solve:
        line["solve"] = solve
    if cpu:
        line["cpu_baseline"] = cpu
    emit(line)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


def time_solve(V, torch, args, g, mg, world, dist, reps=5):
    """ms per MG-PCG solve of the bench grid (all ranks), reference production settings (fem.py:66-70)."""
    f = torch.from_numpy(g.buildLoadVector()).cuda()
    u0 = torch.zeros_like(f); out = torch.empty_like(f)
    kw = dict(mgIterations=1, mgSmoothingIterations=2, fullMultigrid=True, out=out)
    mg.preconditionedConjugateGradient(u0, f, 100, 1e-4, **kw)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(reps):
        mg.preconditionedConjugateGradient(u0, f, 100, 1e-4, **kw)
    ev1.record(); torch.cuda.synchronize()
    t = torch.tensor([ev0.elapsed_time(ev1) / reps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    its = mg.lastIterations
    nn_global = (GRID[0] * world + 1) * (GRID[1] + 1) * (GRID[2] + 1)
    res = {"ms_per_solve": ms, "iterations": int(its), "dofs": 3 * nn_global, "mdof_per_s": 3 * nn_global / ms / 1e3, "n_gpus": world,
           "protocol": "zero initial guess, tol 1e-4, %d coarsening levels, FMG preconditioner, 2 GS sweeps, operator refresh included" % args.levels,
           "residual": mg.lastResidualNorm}
    if world == 1:
        # e2e: densities (host) in -> solve -> compliance + per-voxel gradient (host) out, as VoxelFEMFunction.forward does (fem.py:109-129)
        rho = g.getDensities()
        obj = V.MultigridComplianceObjective(mg)
        obj.tol, obj.zeroInit = 1e-4, True
        obj.updateCache(rho); obj.compliance(); obj.gradient()
        t0 = time.perf_counter()
        for _ in range(3):
            obj.updateCache(rho); J = obj.compliance(); gr = obj.gradient()
        res["e2e_ms_densities_to_gradient"] = (time.perf_counter() - t0) / 3 * 1e3
        res["compliance"] = float(J)
    return res


_JSON_FD = None


def emit(line):
    """the ONE JSON line, on the process's real stdout"""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    # stdout carries the JSON line only: whatever libraries print there (NCCL's version banner, ...) goes to stderr
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--levels", type=int, default=3)
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-solve", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        reference_arm(args, int(os.environ.get("RANK", 0)))
    else:
        our_arm(args)


if __name__ == "__main__":
    main()
