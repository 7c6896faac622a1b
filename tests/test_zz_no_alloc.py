"""vfem_dot_no_alloc (include/vfem_b200.h): the allocation-free dot product (reduction scratch in a static __device__ array,
ndr_b200/csrc/no_alloc.cu) against the existing entry point vfem_mg_dot, through the C ABI, on two calls in a row.
No counterpart in the reference; the file name keeps it last in pytest's collection order (tests/conftest.py)."""
import ctypes as C

import numpy as np
import pytest

from conftest import MATERIAL, bc_path

pytestmark = pytest.mark.gpu

TOL = 1e-13   # the per-kernel parity bar of tests/test_gpu_parity.py, relative to sum |a_i b_i|


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def test_dot_no_alloc_matches_mg_dot_twice():
    import torch
    import ndr_b200.pyVoxelFEM as V
    from ndr_b200._lib import check, lib
    from ndr_b200.voxelfem import DEVICE
    L = lib()
    tps = V.TensorProductSimulator([1, 1, 1], [np.zeros(3), np.array([2.0, 1.0, 1.0])], [48, 24, 24])
    tps.readMaterial(MATERIAL)
    tps.applyDisplacementsAndLoadsFromFile(bc_path("3d/cantilever_flexion.bc"))
    mg = tps.multigridSolver(2)
    n = tps.numDoFs()
    assert n > 2 * 4096 and n % 4096 != 0          # several blocks and a ragged last one
    assert n <= L.vfem_dot_no_alloc_max_n()
    gen = torch.Generator(device="cuda"); gen.manual_seed(5)
    res = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    for call in range(2):                           # two calls in a row share the static scratch
        a = torch.randn(n, dtype=torch.float64, device="cuda", generator=gen)
        b = torch.randn(n, dtype=torch.float64, device="cuda", generator=gen) + 0.25 * a
        torch.cuda.synchronize()                    # torch's stream is done before the library's (legacy default) stream reads
        want = C.c_double()
        check(L.vfem_mg_dot(mg._h, _ptr(a), _ptr(b), DEVICE, C.byref(want)))
        check(L.vfem_dot_no_alloc(_ptr(a), _ptr(b), n, _ptr(res)))
        check(L.vfem_device_synchronize())
        got = float(res.cpu()[0])
        scale = float((a.abs() * b.abs()).sum())
        assert abs(got - want.value) <= TOL * scale, (call, got, want.value)
        assert abs(got - float(torch.dot(a, b))) <= TOL * scale
    # a short vector right after a long one: stale partials of the earlier calls must not leak in
    check(L.vfem_dot_no_alloc(_ptr(a), _ptr(b), 1000, _ptr(res)))
    check(L.vfem_device_synchronize())
    assert abs(float(res.cpu()[0]) - float(torch.dot(a[:1000], b[:1000]))) <= TOL * float((a[:1000].abs() * b[:1000].abs()).sum())
    check(L.vfem_dot_no_alloc(_ptr(a), _ptr(b), 0, _ptr(res)))
    check(L.vfem_device_synchronize())
    assert float(res.cpu()[0]) == 0.0


def test_dot_no_alloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    from ndr_b200._lib import lib
    L = lib()
    a = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = L.vfem_dot_no_alloc(_ptr(a), _ptr(a), L.vfem_dot_no_alloc_max_n() + 1, _ptr(a))
    assert rc != 0
    msg = L.vfem_last_error().decode()
    assert "exceeds the static scratch" in msg and str(L.vfem_dot_no_alloc_max_n()) in msg
