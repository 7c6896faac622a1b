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


def host_threads():
    """physical cores of this box (the reference sizes TBB the same way, training/train_voxelfem.py:38); torchrun exports
    OMP_NUM_THREADS=1, so the oracle's thread count is always set explicitly"""
    try:
        import psutil
        n = psutil.cpu_count(logical=False)
    except Exception:  # noqa: BLE001
        n = None
    return int(n or os.cpu_count() or 1)


def oracle_sim(grid, domain, rho):
    from oracle import oracle as O
    O.build()
    O.set_num_threads(host_threads())
    s = O.OracleSim([0, 0, 0], domain, grid)
    s.read_material(MATERIAL)
    s.set_densities(rho)
    s.apply_bc_file(BC)
    s.set_simp(1.0, 1e-4, 3.0)
    return O, s


def run_cpu_kapply(budget_s, solve=True):
    """CPU restatement of the reference (oracle/): K-apply GDOF/s on the bench grid + one MG-PCG solve (kept for the parity block)."""
    O, s = oracle_sim(GRID, DOMAIN, np.random.default_rng(88).uniform(0, 1, int(np.prod(GRID))))
    u = np.random.default_rng(89).standard_normal((s.nn, 3))
    for _ in range(10):
        s.apply_k(u)
    t0, reps = time.perf_counter(), 0
    while reps < 3 or (time.perf_counter() - t0 < budget_s and reps < 200):
        s.apply_k(u); reps += 1
    dt = (time.perf_counter() - t0) / reps
    out = {"value": 3 * s.nn / dt / 1e9, "unit": "GDOF/s", "cores": O.max_threads(), "kind": "port",
           "sample": "%d K-applies of the 128x64x64 cantilever (seed-88 densities) after 10 warm-up applies, OpenMP oracle restatement; the "
                     "reference binary is unbuildable here (Eigen/TBB/SuiteSparse absent)" % reps, "ms_per_apply": dt * 1e3}
    sol = None
    if solve:
        s.mg_create(3)
        f = s.load_vector()
        t0 = time.perf_counter()
        x, its, hist = s.mg_pcg(np.zeros_like(f), f, 100, 1e-4, 1, 2, True)
        out["solve_ms"] = (time.perf_counter() - t0) * 1e3
        out["solve_iterations"] = int(its)
        sol = {"u": x, "f": f, "iterations": int(its), "residual": float(hist[-1]) if len(hist) else None}
    return out, sol


def reference_arm(args, rank):
    """The reference's CPU algorithm (oracle restatement; the reference binary is unbuildable here) on all physical host
    cores, on the SAME workload as our arm at this --gpus: the (gpus*128)x64x64 grid with the same densities."""
    if rank != 0:
        return
    W, K, world = args.warmup, args.steps, args.gpus
    gE = [GRID[0] * world, GRID[1], GRID[2]]
    dom = [DOMAIN[0] * world, DOMAIN[1], DOMAIN[2]]
    rho = np.random.default_rng(88).uniform(0, 1, int(np.prod(gE))) if world == 1 else layer_densities(0, gE[0], GRID[1] * GRID[2])
    O, s = oracle_sim(gE, dom, rho)
    u = np.random.default_rng(89).standard_normal((s.nn, 3))
    t_pre, pre = time.perf_counter(), 0
    while pre < 50 and time.perf_counter() - t_pre < 20.0:  # steady state (page faults, thread pool, caches) before the W warm-up steps
        s.apply_k(u); pre += 1
    for _ in range(W):
        s.apply_k(u)
    t0 = time.perf_counter()
    for _ in range(K):
        s.apply_k(u)
    dt = (time.perf_counter() - t0) / K
    val = 3 * s.nn / dt / 1e9
    cb = {"value": val, "unit": "GDOF/s", "cores": O.max_threads(), "kind": "port",
          "sample": "%d K-applies (after %d + %d warm-up applies) of the %dx%dx%d cantilever, OpenMP restatement of the reference algorithm "
                    "(reference unbuildable here)" % (K, pre, W, gE[0], gE[1], gE[2])}
    line = {"impl": "reference", "metric": "k_apply_gdof_per_s", "value": val, "unit": "GDOF/s", "n_gpus": args.gpus, "steps": K, "warmup": W,
            "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(gE, world, None, None),
            "cpu_baseline": cb, "e2e": {"value": val, "unit": "GDOF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}
    emit(line)


def workload_config(gE, world, nsets, nn_l):
    cfg = {"workload": "3D cantilever %dx%dx%d Q1 hex (128x64x64 element layers per GPU, slabs along axis 0), matrix-free K(rho)u, "
                       "seed-88 uniform(0,1) densities, E_min=1e-4, p=3" % tuple(gE),
           "dofs": 3 * (gE[0] + 1) * (gE[1] + 1) * (gE[2] + 1)}
    if nsets:
        cfg["l2"] = "ring of %d u/Ku buffer pairs (%.0f MB) > 126 MB L2" % (nsets, nsets * 48 * nn_l / 1e6)
        cfg["halo"] = "none" if world == 1 else "1 node plane (%d B) per interior face per apply, ncclSend/ncclRecv on the solver stream" % ((gE[1] + 1) * (gE[2] + 1) * 24)
    return cfg


def layer_densities(layer_lo, layer_hi, plane_elems):
    """synthetic densities of global element layers [layer_lo, layer_hi): one seeded stream per layer, so that the ghost
    layers of neighbouring ranks hold identical values whatever the decomposition"""
    return np.concatenate([np.random.default_rng([88, k]).uniform(0, 1, plane_elems) for k in range(layer_lo, layer_hi)])


def _ring(tensors):
    import ctypes as C
    return (C.c_void_p * len(tensors))(*[t.data_ptr() for t in tensors])


class Workload:
    """one grid on `world` ranks: simulator (+ slab solver for world > 1), multigrid hierarchy, load vector"""

    def __init__(self, V, S, torch, dist, gE, dom, bc, levels, rank, world, local, rho_mode):
        self.gE, self.world, self.levels = gE, world, levels
        bbmin, bbmax = np.zeros(3), np.array(dom, float)
        plane = gE[1] * gE[2]
        if world == 1:
            tps = V.TensorProductSimulator([1, 1, 1], [bbmin, bbmax], gE)
            tps.readMaterial(MATERIAL)
            tps.setElementDensities(np.random.default_rng(88).uniform(0, 1, tps.numElements()) if rho_mode == "flat" else layer_densities(0, gE[0], plane))
            tps.applyDisplacementsAndLoadsFromFile(bc)
            tps.E_0, tps.E_min, tps.gamma = 1.0, 1e-4, 3.0
            self.mg = tps.multigridSolver(levels)
            self.mg_halo = None
        else:
            comm = S.make_nccl_comm(dist, rank, world)
            sl = S.SlabSolver(gE, bbmin, bbmax, levels, rank, world, comm, S.dist_host_sum(dist, torch.device("cuda", local)), MATERIAL)
            tps, lay = sl.tps, sl.layout
            tps.setElementDensities(layer_densities(lay.layer_lo, lay.layer_hi, lay.plane_elems))
            sl.apply_bc_file(bc)
            tps.E_0, tps.E_min, tps.gamma = 1.0, 1e-4, 3.0
            self.mg = sl.multigrid()
            self.mg_halo = self.mg._h
            self.slab = sl
        self.tps = tps
        self.nn_l, self.ne_l = tps.numNodes(), tps.numElements()
        self.nn_global = (gE[0] + 1) * (gE[1] + 1) * (gE[2] + 1)
        self.ne_global = gE[0] * gE[1] * gE[2]


def time_kapply(L, torch, dist, wl, W, K, seed, preheat_s=0.3):
    """K matrix-free applies (N > 1: + the one-plane halo shift per interior face, hidden behind the interior layers) issued by ONE
    C-ABI call, so the Python launch rate is outside the timed region; CUDA events on the library's stream (the legacy default
    stream, which is also torch's current stream here); max over ranks.  Returns (ms total over K steps, ms per kernel alone, launches)."""
    world = wl.world
    nsets = max(2, int(np.ceil(2.5 * L2_BYTES / (16 * 3 * wl.nn_l))))  # ring of u/Ku pairs larger than L2
    if world > 1:   # the slabs differ in size (ghost layers): every rank must issue the SAME number of collective applies
        t = torch.tensor([nsets], dtype=torch.int64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        nsets = int(t.item())
    gen = torch.Generator(device="cuda"); gen.manual_seed(seed)
    us = [torch.randn(wl.nn_l, 3, dtype=torch.float64, device="cuda", generator=gen) for _ in range(nsets)]
    outs = [torch.zeros(wl.nn_l, 3, dtype=torch.float64, device="cuda") for _ in range(nsets)]
    pu, po = _ring(us), _ring(outs)

    def run(start, count, halo=True):
        if wl.mg_halo and halo:
            _lib_check(L.vfem_mg_apply_k_halo_ring(wl.mg_halo, pu, po, nsets, start, count))
        else:
            _lib_check(L.vfem_tps_apply_k_ring(wl.tps._h, pu, po, nsets, start, count))

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    if os.environ.get("VFEM_BENCH_TRACE"):
        import faulthandler
        faulthandler.dump_traceback_later(45, exit=False)   # where is a rank that never leaves the preheat?

    def trace(msg):
        if os.environ.get("VFEM_BENCH_TRACE"):
            print("[rank %s %.2f] time_kapply %s" % (os.environ.get("RANK", "0"), time.time() % 1000, msg), file=sys.stderr, flush=True)

    # untimed: every ring buffer touched, clocks ramped to their load state, then the W warm-up steps
    # (a fixed count derived from the GLOBAL size: the ranks must issue the same number of collective applies)
    preheat_s = float(os.environ.get("VFEM_BENCH_PREHEAT_S", preheat_s))   # profiling runs (ncu launch lists) shorten the untimed ramp
    reps = max(1, min(500 if world == 1 else 40, int(preheat_s / (nsets * 3.0 * wl.nn_global / world / 60e9))))
    for _ in range(reps):
        run(0, nsets)
        if world > 1:
            torch.cuda.synchronize()
    trace("preheat done (%d x %d applies)" % (reps, nsets))
    run(0, W)
    sync_all()
    trace("warm-up done")
    launches0 = L.vfem_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    t_issue = time.perf_counter()
    run(W, K)
    t_issue = (time.perf_counter() - t_issue) * 1e3 / K   # host time to ENQUEUE one step (N > 1: includes the in-flight bound's waits)
    ev1.record()
    sync_all()
    trace("timed region done; host issue %.1f us per step" % (t_issue * 1e3))
    ms = ev0.elapsed_time(ev1)
    launches = L.vfem_launch_count() - launches0
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    # the dominant kernel alone (local apply, no halo exchange): the roofline leg
    torch.cuda.synchronize()
    ev0.record()
    run(W, K, halo=False)
    ev1.record(); torch.cuda.synchronize()
    kms = ev0.elapsed_time(ev1) / K
    return ms, kms, int(launches), nsets, (us, outs)


MARCH_SHAPES = [(8, 64), (16, 32), (15, 33), (7, 65), (6, 64), (4, 64), (8, 32), (8, 33), (5, 64), (4, 33), (6, 65), (5, 33), (11, 33)]


def kernel_label(tps):
    info = tps.applyInfo()
    if not info["march"]:
        return "applyFineKernel<3,0> (per-node gather)"
    ry, zt = MARCH_SHAPES[info["tile_shape"]] if info["tile_shape"] < len(MARCH_SHAPES) else (0, 0)
    name = "applyMarchKernel" if os.environ.get("VFEM_APPLY_VARIANT") == "4" else "applyMarchTmaKernel"   # sim.cu: variant 4 = first generation
    return name + "<MODE=0, SPARSE=%s, RY=%d, ZT=%d> grid (%d,%d,%d), %d layers per chunk" % (
        "true" if info["k0_structure"] == 2 else "false", ry, zt, info["grid_x"], info["grid_y"], info["grid_z"], info["layers_per_chunk"])


def roofline_block(wl, kms, hbm_peak, peak_src):
    set_bytes = algorithmic_bytes(wl.nn_l, wl.ne_l)
    achieved = set_bytes / (kms * 1e-3) / 1e9
    r = {"bound": "hbm", "kernel": kernel_label(wl.tps), "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
         "traffic": None, "peak_source": peak_src, "algorithmic_bytes_per_launch": set_bytes, "kernel_ms": kms,
         "gdof_per_s_per_gpu": 3.0 * wl.nn_l / (kms * 1e-3) / 1e9}
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):  # dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full capture, keyed by grid and kernel source
        t = json.load(open(tp))
        key = "%dx%dx%d" % tuple(wl.gE)
        ent = t.get("apply_march", {}).get(key)
        if ent and ent.get("kernel_source_sha16") == kernel_source_sha():
            r["traffic"] = ent["dram_bytes_per_launch"]
            r["traffic_source"] = ent.get("source")
    return r


def kernel_source_sha():
    import hashlib
    h = hashlib.sha256()
    for f in ("apply_march.cuh", "apply_march_tma.cuh"):
        h.update(open(os.path.join(ROOT, "ndr_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def _lib_check(rc):
    from ndr_b200 import _lib
    _lib.check(rc)


def our_arm(args):
    import torch
    import torch.distributed as dist
    from ndr_b200 import _lib
    from ndr_b200 import slab as S
    import ndr_b200.pyVoxelFEM as V

    rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1)); local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    _lib.check(_lib.lib().vfem_set_device(local))
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=180))
        # a collective that never completes must not hold the box for the driver's whole limit
        import signal
        def _alarm(signum, frame):
            sys.stderr.write("bench.py: rank %d made no progress for 900 s - aborting\n" % rank)
            os._exit(3)
        signal.signal(signal.SIGALRM, _alarm)
        signal.alarm(900)
    W, K = args.warmup, args.steps
    hbm_peak, peak_src = measured_peaks()
    L = _lib.lib()
    t_start = time.perf_counter()

    # ---- headline workload: weak scaling, one 128x64x64 block of element layers per GPU, slabs along axis 0 ----
    gE = [GRID[0] * world, GRID[1], GRID[2]]
    # coarsening levels: the reference's production value 3 (fem.py --mgl 3); the weak-scaling grid grows along axis 0, and with
    # it the plane chain of the coarsest direct solve, so the hierarchy is deepened until the coarsest grid has <= 64 layers
    # (N = 8: 4 levels; measured 46.0 ms per solve against 66.9 ms with 3, same 10 iterations - profiles/r02c_weak_levels_n8.txt)
    levels = args.levels
    if levels <= 0:
        levels = 3
        while (gE[0] >> levels) > 64 and (GRID[0] % (1 << (levels + 1))) == 0 and min(GRID[1], GRID[2]) % (1 << (levels + 1)) == 0:
            levels += 1
    wl = Workload(V, S, torch, dist, gE, [DOMAIN[0] * world, DOMAIN[1], DOMAIN[2]], BC, levels, rank, world, local, "flat")
    sampler = ClockSampler(local); sampler.start()
    ms, kms, launches, nsets, bufs = time_kapply(L, torch, dist, wl, W, K, 89 + rank)
    clocks = sampler.stop()
    value = 3.0 * wl.nn_global * K / (ms * 1e-3) / 1e9
    roofline = roofline_block(wl, kms, hbm_peak, peak_src)

    # ---- MG-PCG solve over all ranks (collective: every rank takes part) ----
    solve = gpu_sol = None
    if not args.no_solve:
        solve, gpu_sol = time_solve(V, torch, wl, dist, keep_solution=(world == 1))

    # ---- e2e: the reference-facing call with HOST buffers, H2D + D2H inside the timed region; every rank runs its slab's
    # host-array apply at the same time (they share the host memory system and PCIe switches), time = max over ranks ----
    e2e = time_e2e(torch, dist, wl, world, K)
    del bufs
    torch.cuda.empty_cache()

    # ---- the other BASELINE configurations (strong scaling across the N ranks) ----
    configs = {}
    if not args.no_extra:
        configs = extra_configs(args, V, S, torch, dist, L, rank, world, local, hbm_peak, peak_src)

    if rank != 0:
        if world > 1:
            dist.barrier(); dist.destroy_process_group()
        return

    cpu = parity = None
    if world == 1 and not args.no_cpu:
        cpu, cpu_sol = run_cpu_kapply(args.cpu_seconds, solve=not args.no_solve)
        if cpu_sol is not None and gpu_sol is not None:
            parity = parity_block(gpu_sol, cpu_sol)
    line = {"metric": "k_apply_gdof_per_s", "value": value, "unit": "GDOF/s", "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(gE, world, nsets, wl.nn_l),
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline}
    line["config"]["timing"] = "the K applies are issued by one C-ABI call (vfem_tps_apply_k_ring / vfem_mg_apply_k_halo_ring) between two CUDA events"
    if solve:
        line["solve"] = solve
    if cpu:
        line["cpu_baseline"] = cpu
    if parity:
        line["parity"] = parity
    if configs:
        line["configs"] = configs
    line["bench_wall_s"] = time.perf_counter() - t_start
    emit(line)
    if world > 1:
        dist.barrier(); dist.destroy_process_group()


def parity_block(gpu_sol, cpu_sol):
    """GPU MG-PCG solve vs the oracle's solve of the same system (both: zero start, tol 1e-4, FMG, 2 sweeps)"""
    ug, uo, f = gpu_sol["u"], cpu_sol["u"], cpu_sol["f"]
    cg, co = 0.5 * float(np.sum(f * ug)), 0.5 * float(np.sum(f * uo))
    return {"against": "oracle restatement, same grid / loads / densities, same solver settings",
            "u_rel_err_max": float(np.max(np.abs(ug - uo)) / np.max(np.abs(uo))), "compliance_rel_err": abs(cg - co) / abs(co),
            "iterations_gpu": gpu_sol["iterations"], "iterations_oracle": cpu_sol["iterations"],
            "residual_gpu": gpu_sol["residual"], "residual_oracle": cpu_sol["residual"], "compliance_gpu": cg, "compliance_oracle": co}


def time_e2e(torch, dist, wl, world, K):
    tps, nn_l = wl.tps, wl.nn_l
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ke = max(3, min(K, 20))
    res = {}
    for kind in ("pinned", "pageable"):
        if kind == "pinned":
            uh = torch.randn(nn_l, 3, dtype=torch.float64).pin_memory().numpy()
            oh = torch.empty(nn_l, 3, dtype=torch.float64).pin_memory().numpy()
        else:  # what fem.py's callers pass: plain numpy arrays
            uh = np.random.default_rng(5).standard_normal((nn_l, 3)); oh = np.empty((nn_l, 3))
        for _ in range(3):
            tps.applyK(uh, out=oh)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter(); ev0.record()
        for _ in range(ke):
            tps.applyK(uh, out=oh)
        ev1.record(); torch.cuda.synchronize()
        wall = (time.perf_counter() - t0) / ke * 1e3
        t = torch.tensor([max(ev0.elapsed_time(ev1) / ke, wall)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res[kind] = float(t.item())
    h2d = int(nn_l * 24)
    return {"value": 3.0 * wl.nn_global / (res["pinned"] * 1e-3) / 1e9, "unit": "GDOF/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": h2d,
            "ms_per_step": res["pinned"], "buffers": "pinned host arrays",
            "pageable": {"value": 3.0 * wl.nn_global / (res["pageable"] * 1e-3) / 1e9, "ms_per_step": res["pageable"]},
            "call": "pyVoxelFEM TensorProductSimulator.applyK(numpy) -> vfem_tps_apply_k(VFEM_HOST); all %d ranks run their slab's call "
                    "concurrently, time = max over ranks (host clock around the loop, device synchronised)" % world}


def time_solve(V, torch, wl, dist, reps=5, keep_solution=False, e2e=True):
    """ms per MG-PCG solve (all ranks), reference production settings (fem.py:66-70)."""
    g, mg, world = wl.tps, wl.mg, wl.world
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
    res = {"ms_per_solve": ms, "iterations": int(its), "dofs": 3 * wl.nn_global, "mdof_per_s": 3 * wl.nn_global / ms / 1e3, "n_gpus": world,
           "ms_per_iteration": ms / max(1, int(its)),
           "protocol": "zero initial guess, tol 1e-4, %d coarsening levels, FMG preconditioner, 2 GS sweeps, operator refresh included" % wl.levels,
           "residual": mg.lastResidualNorm}
    sol = None
    if keep_solution:
        sol = {"u": out.cpu().numpy(), "iterations": int(its), "residual": float(mg.lastResidualNorm)}
    if world == 1 and e2e:
        # e2e: densities (host) in -> solve -> compliance + per-voxel gradient (host) out, as VoxelFEMFunction.forward does (fem.py:109-129)
        rho = g.getDensities()
        obj = V.MultigridComplianceObjective(mg)
        obj.tol, obj.zeroInit = 1e-4, True
        for _ in range(2):
            obj.updateCache(rho); obj.compliance(); obj.gradient()
        ts = []
        for _ in range(5):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            obj.updateCache(rho); J = obj.compliance(); gr = obj.gradient()
            ts.append((time.perf_counter() - t0) * 1e3)
        res["e2e_ms_densities_to_gradient"] = float(np.median(ts))
        res["e2e_ms_densities_to_gradient_min_max"] = [float(min(ts)), float(max(ts))]
        res["compliance"] = float(J)
    return res, sol


def extra_configs(args, V, S, torch, dist, L, rank, world, local, hbm_peak, peak_src):
    """BASELINE configs[2] and [3] in the same driver-run record: K-apply (+ halo) and one MG-PCG solve, STRONG scaling over the
    `world` ranks.  256x128x128 uses the reference's own mgl = 3 (logs/slurm/gt/c1001.log); 512x256x256 needs 5 levels (its
    mgl-3 coarsest grid, 64x32x32, is too large for the block-tridiagonal factor)."""
    fix3 = os.path.join(FIX, "bcs", "3d")
    cases = [("bridge_256x128x128", [256, 128, 128], [4.0, 2.0, 1.0], os.path.join(fix3, "bridge.bc"), 3),
             ("cantilever_512x256x256", [512, 256, 256], [2.0, 1.0, 1.0], os.path.join(fix3, "cantilever_flexion.bc"), 5)]
    out = {}
    for name, gE, dom, bc, levels in cases:
        if (gE[0] // world) % (1 << levels) or world > 8:
            continue
        if name.startswith("cantilever_512") and world not in (1, 8) and not args.all_extra:
            continue
        try:
            wl = Workload(V, S, torch, dist, gE, dom, bc, levels, rank, world, local, "layers")
            ms, kms, launches, nsets, bufs = time_kapply(L, torch, dist, wl, 3, 10, 189 + rank, preheat_s=0.1)
            del bufs
            solve, _ = time_solve(V, torch, wl, dist, reps=2, e2e=False)
            ent = {"grid": gE, "dofs": 3 * wl.nn_global, "n_gpus": world, "scaling": "strong", "levels": levels,
                   "k_apply_gdof_per_s": 3.0 * wl.nn_global * 10 / (ms * 1e-3) / 1e9, "k_apply_ms": ms / 10,
                   "roofline": roofline_block(wl, kms, hbm_peak, peak_src), "solve": solve}
            out[name] = ent
            del wl
            torch.cuda.empty_cache()
        except Exception as e:  # noqa: BLE001  (a config that does not fit must not take the headline down)
            out[name] = {"error": "%s: %s" % (type(e).__name__, e)}
            if world > 1:
                raise
    if world == 1:  # BASELINE configs[4]: one neural-design step (MLP forward, VoxelFEM solve + gradient, backward, Adam) at 256x128x128
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import neural_loop
            out["neural_loop_256x128x128"] = neural_loop.run([256, 128, 128], steps=3)
            torch.cuda.empty_cache()
        except Exception as e:  # noqa: BLE001
            out["neural_loop_256x128x128"] = {"error": "%s: %s" % (type(e).__name__, e)}
    return out


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
    ap.add_argument("--levels", type=int, default=0, help="coarsening levels of the headline solve (0 = 3, deepened for long weak-scaling grids)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-solve", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip the 256x128x128 / 512x256x256 configurations")
    ap.add_argument("--all-extra", action="store_true", help="512x256x256 at every N (default: N = 1 and 8)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)
    if args.impl == "reference":
        reference_arm(args, int(os.environ.get("RANK", 0)))
    else:
        our_arm(args)


if __name__ == "__main__":
    main()
