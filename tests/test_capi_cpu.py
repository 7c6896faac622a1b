"""CPU-side checks of the drop-in boundary: the C-ABI library builds, loads and exports every symbol that
include/vfem_b200.h declares; the product never imports the oracle."""
import ctypes
import os
import re

from conftest import ROOT


def test_library_exports_every_declared_symbol():
    from ndr_b200 import _lib, build
    so = build.build()
    protos = _lib.parse_header()
    assert len(protos) >= 70
    h = ctypes.CDLL(so)
    missing = [n for n in protos if not hasattr(h, n)]
    assert not missing, missing
    # every vfem_* name mentioned in the header is covered by the parser
    txt = re.sub(r"/\*.*?\*/", "", open(_lib.HEADER).read(), flags=re.S)
    names = set(re.findall(r"\b(vfem_\w+)\s*\(", txt))
    assert names <= set(protos), names - set(protos)


def test_no_device_reports_error_not_fallback():
    """Without a GPU the factory must fail loudly (no CPU fallback)."""
    import numpy as np
    from ndr_b200 import _lib
    L = _lib.lib()
    cnt = ctypes.c_int(0)
    if L.vfem_device_count(ctypes.byref(cnt)) == 0 and cnt.value > 0:
        return  # a GPU is present; covered by the gpu tests
    deg = np.ones(3, np.uint64); el = np.array([4, 4, 4], np.uint64)
    lo, hi = np.zeros(3), np.ones(3)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    h = ctypes.c_void_p()
    rc = L.vfem_tps_create(3, p(deg), p(lo), p(hi), p(el), ctypes.byref(h))
    assert rc != 0 and b"no CUDA device" in L.vfem_last_error()


def test_product_does_not_touch_oracle():
    pkg = os.path.join(ROOT, "ndr_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inl")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("oracle-free", ""), os.path.join(dirpath, f)


def test_bc_parser_matches_oracle_parser(oracle_mod):
    import numpy as np
    from ndr_b200.voxelfem import parse_bc_regions
    from conftest import bc_path
    for name, dim, dom in (("3d/cantilever_flexion.bc", 3, [2, 1, 1]), ("3d/bridge.bc", 3, [4, 2, 1]), ("2d/mbb_beam.bc", 2, [3, 1]),
                           ("2d/bridge.bc", 2, [2, 1]), ("2d/roof.bc", 2, [2, 1]), ("2d/short_beam.bc", 2, [2, 1])):
        a = parse_bc_regions(bc_path(name), np.zeros(dim), dom, dim)
        b = oracle_mod.parse_bc_file(bc_path(name), np.zeros(dim), dom, dim)
        assert len(a) == len(b)
        for ra, rb in zip(a, b):
            assert ra[0] == rb[0]
            for x, y in zip(ra[1:], rb[1:]):
                np.testing.assert_array_equal(x, y)
