#!/usr/bin/env python
"""Times applyMarchKernel for forced tile shapes / chunk lengths (VFEM_MARCH_SHAPE, VFEM_MARCH_XC) on a GPU box.
Usage: python tools/march_sweep.py [EX EY EZ]   -> one line per configuration: kernel us, GB/s, fraction of MEASURED_PEAKS hbm."""
import json, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import ndr_b200.pyVoxelFEM as V

grid = [int(a) for a in sys.argv[1:4]] if len(sys.argv) >= 4 else [128, 64, 64]
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
MAT = os.path.join(ROOT, "tests", "golden", "fixtures", "materials", "B9Creator.material")
shapes = {0: (8, 64), 1: (16, 32), 2: (15, 33), 3: (7, 65), 4: (6, 64), 5: (4, 64), 6: (8, 32), 7: (8, 33), 8: (5, 64), 9: (4, 33), 10: (6, 65), 11: (5, 33), 12: (11, 33)}
cases = [(None, None)] + [(s, None) for s in shapes] + [(0, x) for x in (5, 9, 17, 33)] + [(4, x) for x in (7, 9, 13)] + [(5, x) for x in (5, 9, 13)]
if os.environ.get("SWEEP_SHORT"): cases = [(0, None), (4, None), (5, None), (3, None)]
if os.environ.get("SWEEP_CASES"): cases = [tuple(None if v == "-" else int(v) for v in c.split(":")) for c in os.environ["SWEEP_CASES"].split(",")]
for s, xc in cases:
    for k, v in (("VFEM_MARCH_SHAPE", s), ("VFEM_MARCH_XC", xc)):
        if v is None: os.environ.pop(k, None)
        else: os.environ[k] = str(v)
    g = V.TensorProductSimulator([1, 1, 1], [np.zeros(3), np.array([2.0, 1.0, 1.0])], grid)
    g.readMaterial(MAT)
    g.setElementDensities(np.random.default_rng(88).uniform(0, 1, g.numElements()))
    g.E_0, g.E_min, g.gamma = 1.0, 1e-4, 3.0
    nn, ne = g.numNodes(), g.numElements()
    nsets = max(2, int(np.ceil(2.5 * 126e6 / (48 * nn))))
    us = [torch.randn(nn, 3, dtype=torch.float64, device="cuda") for _ in range(nsets)]
    outs = [torch.empty(nn, 3, dtype=torch.float64, device="cuda") for _ in range(nsets)]
    for k in range(10): g.applyK(us[k % nsets], out=outs[k % nsets])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    K = 200 if nn < 2e6 else 50
    e0.record()
    for k in range(K): g.applyK(us[k % nsets], out=outs[k % nsets])
    e1.record(); torch.cuda.synchronize()
    us_ = e0.elapsed_time(e1) / K * 1e3
    gbs = 8 * (6 * nn + ne) / (us_ * 1e-6) / 1e9
    print("shape %-8s XC %-4s info %s : %8.2f us  %7.1f GB/s  frac %.3f  %6.1f GDOF/s" % (shapes.get(s, "auto"), xc, g.applyInfo(), us_, gbs, gbs / peak, 3 * nn / us_ / 1e3), flush=True)
    del g, us, outs
