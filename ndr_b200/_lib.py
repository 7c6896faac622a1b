"""ctypes loader for libvfem_b200.so.

Prototypes are read from include/vfem_b200.h so the Python side cannot drift from the C ABI.
There is no CPU fallback: if the CUDA library is missing or cannot be loaded, importing fails.
"""
import ctypes as C
import os
import re

_HERE = os.path.dirname(os.path.abspath(__file__))
HEADER = os.path.join(_HERE, "..", "include", "vfem_b200.h")
SO_PATH = os.path.join(_HERE, "libvfem_b200.so")

SOLVE_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_size_t, C.c_void_p)
PCG_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_double)
RESIDUAL_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_size_t, C.c_double)
FILTER_CB = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t)

_SCALARS = {"int": C.c_int, "size_t": C.c_size_t, "double": C.c_double, "unsigned long long": C.c_ulonglong}
_HANDLES = {"vfem_tps_t", "vfem_mg_t", "vfem_obj_t", "vfem_top_t", "vfem_comm_t"}
_CALLBACKS = {"vfem_solve_cb": SOLVE_CB, "vfem_pcg_cb": PCG_CB, "vfem_residual_cb": RESIDUAL_CB}


def _ctype(decl):
    decl = decl.strip()
    if decl == "void":
        return None
    base = re.sub(r"\b\w+$", "", decl).strip() if not decl.endswith("*") else decl  # drop the parameter name
    base = base.replace("const ", "").replace(" const", "").strip()
    if base.endswith("*"):
        return C.c_void_p  # every array / out-parameter is passed as a raw pointer
    if base in _HANDLES:
        return C.c_void_p
    if base in _CALLBACKS:
        return _CALLBACKS[base]
    return _SCALARS[base]


def parse_header(path=HEADER):
    """Returns {name: (restype, [argtypes])} for every function declared in the header."""
    txt = open(path).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    protos = {}
    for m in re.finditer(r"^\s*(const char \*|void|int|size_t|unsigned long long)\s*(vfem_\w+)\s*\(([^;{]*?)\)\s*;", txt, flags=re.M | re.S):
        ret, name, args = m.group(1).strip(), m.group(2), " ".join(m.group(3).split())
        restype = C.c_char_p if ret.startswith("const char") else (None if ret == "void" else _SCALARS[ret])
        argtypes = [] if args in ("void", "") else [_ctype(a) for a in args.split(",")]
        protos[name] = (restype, argtypes)
    return protos


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError("ndr_b200: %s is missing - run `python ndr_b200/build.py` (nvcc, sm_100a). "
                              "There is no CPU fallback." % SO_PATH)
        handle = C.CDLL(SO_PATH, mode=C.RTLD_GLOBAL)
        for name, (restype, argtypes) in parse_header().items():
            fn = getattr(handle, name)
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = handle
    return _lib


def check(rc):
    if rc != 0:
        raise RuntimeError(lib().vfem_last_error().decode())
