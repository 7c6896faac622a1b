"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes front-end of the CPU restatement of the reference algorithm (oracle/*.hpp).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module, and only as the checker / CPU baseline.  The product
(ndr_b200/) never imports it.

The .bc / .material parsing here restates MeshFEM BoundaryConditions.cc:218-388 and
Materials.cc (isotropic_material) for the subset the shipped fixtures use.
"""
import ctypes as C
import json
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

c_dp = C.POINTER(C.c_double)


def build(force=False):
    so = os.path.join(_HERE, "libvfem_oracle.so")
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".hpp", ".cpp"))]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "libvfem_oracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "libvfem_oracle.so")
        if not os.path.exists(so):
            build()
        _LIB = C.CDLL(so)
        _LIB.orc_last_error.restype = C.c_char_p
        for name in ("orc_num_nodes", "orc_num_elements", "orc_mg_level_nodes", "orc_mg_level_elements", "orc_top_last_iters"):
            getattr(_LIB, name).restype = C.c_size_t
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _chk(rc):
    if rc != 0:
        raise RuntimeError(lib().orc_last_error().decode())


def parse_bc_file(path, bbmin, bbmax, dim):
    """Regions of a .bc file as (kind, lo, hi, mask, value) tuples (BoundaryConditions.cc:251-388)."""
    with open(path) as f:
        params = json.load(f)
    bbmin = np.asarray(bbmin, float)
    bbmax = np.asarray(bbmax, float)
    out = []
    for r in params["regions"]:
        typ = r["type"]
        mask = np.ones(dim, np.uint8)
        if typ.startswith("dirichlet"):
            rest = typ[9:]
            ln = 0
            for ch in rest:
                if ch < "x" or ch > "z":
                    break
                ln += 1
            if ln > 3:
                raise RuntimeError("invalid mask")
            if ln > 0:
                mask = np.zeros(dim, np.uint8)
                for ch in rest[:ln]:
                    k = "xyz".index(ch)
                    if k < dim:
                        mask[k] = 1
            typ = "dirichlet" + rest[ln:]
        if "box" in r:
            lo = np.array(r["box"]["minCorner"], float)[:dim]
            hi = np.array(r["box"]["maxCorner"], float)[:dim]
        elif "box%" in r:
            lo = np.array(r["box%"]["minCorner"], float)[:dim]
            hi = np.array(r["box%"]["maxCorner"], float)[:dim]
            lo = bbmin + lo * (bbmax - bbmin)  # Geometry.hh:259-262
            hi = bbmin + hi * (bbmax - bbmin)
        else:
            raise RuntimeError("unsupported region geometry")
        val = np.array(r["value"], float)[:dim]
        if typ == "force":
            out.append(("force", lo, hi, mask, val))
        elif typ == "dirichlet":
            out.append(("dirichlet", lo, hi, mask, val))
        else:
            raise RuntimeError("Invalid type '%s'" % typ)
    return out


class OracleSim:
    """TensorProductSimulator + MultigridSolver + TopologyOptimizationProblem of the oracle."""

    def __init__(self, bbmin, bbmax, elems):
        self.dim = len(elems)
        self.bbmin = _f64(bbmin)
        self.bbmax = _f64(bbmax)
        self.elems = np.ascontiguousarray(elems, dtype=np.uint64)
        self.h = C.c_void_p()
        _chk(lib().orc_sim_create(self.dim, _p(self.bbmin), _p(self.bbmax), _p(self.elems), C.byref(self.h)))
        self.nn = lib().orc_num_nodes(self.h)
        self.ne = lib().orc_num_elements(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            lib().orc_sim_destroy(self.h)
            self.h = None

    # --- setup ---
    def set_isotropic(self, E, nu):
        _chk(lib().orc_set_isotropic(self.h, C.c_double(E), C.c_double(nu)))

    def read_material(self, path):
        with open(path) as f:
            m = json.load(f)
        if m["type"] != "isotropic_material":
            raise RuntimeError("unsupported material type")
        self.set_isotropic(m["young"], m["poisson"])

    def set_simp(self, E0, Emin, gamma):
        _chk(lib().orc_set_simp(self.h, C.c_double(E0), C.c_double(Emin), C.c_double(gamma)))

    def apply_bc_file(self, path):
        for kind, lo, hi, mask, val in parse_bc_file(path, self.bbmin, self.bbmax, self.dim):
            lo, hi, val = _f64(lo), _f64(hi), _f64(val)
            if kind == "force":
                _chk(lib().orc_add_force(self.h, _p(lo), _p(hi), _p(val)))
            else:
                mask = np.ascontiguousarray(mask, np.uint8)
                _chk(lib().orc_add_dirichlet(self.h, _p(lo), _p(hi), _p(mask), _p(val)))

    def set_densities(self, rho):
        rho = _f64(rho).ravel()
        assert rho.size == self.ne
        _chk(lib().orc_set_densities(self.h, _p(rho)))

    def densities(self):
        out = np.empty(self.ne)
        _chk(lib().orc_get_densities(self.h, _p(out)))
        return out

    def K0(self):
        ks = self.dim * 2 ** self.dim
        out = np.empty((ks, ks))
        _chk(lib().orc_get_k0(self.h, _p(out)))
        return out

    def load_vector(self):
        out = np.empty((self.nn, self.dim))
        _chk(lib().orc_load_vector(self.h, _p(out)))
        return out

    def node_position(self, ni):
        out = np.empty(self.dim)
        _chk(lib().orc_node_position(self.h, C.c_size_t(ni), _p(out)))
        return out

    # --- operators ---
    def apply_k(self, u):
        u = _f64(u)
        out = np.empty((self.nn, self.dim))
        _chk(lib().orc_apply_k(self.h, _p(u), _p(out)))
        return out

    def solve_direct(self, f):
        f = _f64(f)
        out = np.empty((self.nn, self.dim))
        _chk(lib().orc_solve_direct(self.h, _p(f), _p(out)))
        return out

    def compliance_gradient(self, u):
        u = _f64(u)
        out = np.empty(self.ne)
        _chk(lib().orc_compliance_gradient(self.h, _p(u), _p(out)))
        return out

    # --- multigrid ---
    def mg_create(self, levels):
        _chk(lib().orc_mg_create(self.h, C.c_size_t(levels)))
        self.levels = levels

    def mg_set_flags(self, symmetric_gs=True, build_block=True, build_finest=False):
        _chk(lib().orc_mg_set_flags(self.h, int(symmetric_gs), int(build_block), int(build_finest)))

    def mg_level_nodes(self, l):
        return lib().orc_mg_level_nodes(self.h, C.c_size_t(l))

    def mg_level_elements(self, l):
        return lib().orc_mg_level_elements(self.h, C.c_size_t(l))

    def mg_level_dmask(self, l):
        out = np.empty((self.mg_level_nodes(l), self.dim), np.uint8)
        _chk(lib().orc_mg_level_dmask(self.h, C.c_size_t(l), _p(out)))
        return out

    def mg_update_stiffness(self):
        _chk(lib().orc_mg_update_stiffness(self.h))

    def mg_update_blockks(self):
        _chk(lib().orc_mg_update_blockks(self.h))

    def mg_level_ke(self, l, e):
        ks = self.dim * 2 ** self.dim
        out = np.empty((ks, ks))
        _chk(lib().orc_mg_level_ke(self.h, C.c_size_t(l), C.c_size_t(e), _p(out)))
        return out

    def mg_level_stencil(self, l):
        nst = 3 ** self.dim
        out = np.empty((self.mg_level_nodes(l), nst, self.dim, self.dim))
        _chk(lib().orc_mg_level_stencil(self.h, C.c_size_t(l), _p(out)))
        return out

    def mg_apply_k(self, l, u):
        u = _f64(u)
        out = np.empty_like(u)
        _chk(lib().orc_mg_apply_k(self.h, C.c_size_t(l), _p(u), _p(out)))
        return out

    def mg_residual(self, l, u, b):
        u, b = _f64(u), _f64(b)
        out = np.empty_like(u)
        _chk(lib().orc_mg_residual(self.h, C.c_size_t(l), _p(u), _p(b), _p(out)))
        return out

    def mg_smooth(self, l, u, b, forward=True, serial=False):
        u, b = _f64(u).copy(), _f64(b)
        fn = lib().orc_mg_smooth_serial if serial else lib().orc_mg_smooth_colored
        _chk(fn(self.h, C.c_size_t(l), _p(u), _p(b), int(forward)))
        return u

    def mg_restrict(self, lf, vf):
        vf = _f64(vf)
        out = np.empty((self.mg_level_nodes(lf + 1), self.dim))
        _chk(lib().orc_mg_restrict(self.h, C.c_size_t(lf), _p(vf), _p(out)))
        return out

    def mg_interpolate(self, lf, vc, accum_into=None):
        vc = _f64(vc)
        if accum_into is None:
            out = np.empty((self.mg_level_nodes(lf), self.dim))
            _chk(lib().orc_mg_interpolate(self.h, C.c_size_t(lf), _p(vc), _p(out), 0))
        else:
            out = _f64(accum_into).copy()
            _chk(lib().orc_mg_interpolate(self.h, C.c_size_t(lf), _p(vc), _p(out), 1))
        return out

    def mg_zero_dirichlet(self, l, u):
        u = _f64(u).copy()
        _chk(lib().orc_mg_zero_dirichlet(self.h, C.c_size_t(l), _p(u)))
        return u

    def mg_color_order(self):
        out = np.empty(self.nn, np.int32)
        _chk(lib().orc_mg_color_order(self.h, _p(out)))
        return out

    def mg_solve(self, u, f, num_steps, nu, stiffness_updated=False, zero_dirichlet=False, fmg=False):
        u, f = _f64(u), _f64(f)
        out = np.empty_like(u)
        _chk(lib().orc_mg_solve(self.h, _p(u), _p(f), C.c_size_t(num_steps), C.c_size_t(nu), int(stiffness_updated),
                                int(zero_dirichlet), int(fmg), _p(out)))
        return out

    def mg_pcg(self, u, b, max_iter, tol, mg_iterations=1, nu=1, fmg=False):
        """Returns (x, iterations, residual-norm history)."""
        u, b = _f64(u), _f64(b)
        out = np.empty_like(u)
        iters = C.c_size_t(0)
        hist = np.zeros(max(max_iter, 1))
        _chk(lib().orc_mg_pcg(self.h, _p(u), _p(b), C.c_size_t(max_iter), C.c_double(tol), C.c_size_t(mg_iterations),
                              C.c_size_t(nu), int(fmg), _p(out), C.byref(iters), _p(hist), C.c_size_t(hist.size)))
        return out, iters.value, hist[: iters.value]

    # --- topology optimisation ---
    def top_create(self, use_mg, volfrac, filters=()):
        """filters: sequence of ("smoothing", radius) / ("projection", beta)."""
        ft = np.array([0 if f[0] == "smoothing" else 1 for f in filters], np.int32)
        fp = np.array([float(f[1]) for f in filters], np.float64)
        _chk(lib().orc_top_create(self.h, int(use_mg), C.c_double(volfrac), len(filters), _p(ft), _p(fp)))

    def top_set_solver(self, cg_iter=100, tol=1e-5, mg_iterations=1, nu=2, fmg=True, zero_init=False):
        _chk(lib().orc_top_set_solver(self.h, C.c_size_t(cg_iter), C.c_double(tol), C.c_size_t(mg_iterations), C.c_size_t(nu),
                                      int(fmg), int(zero_init)))

    def top_set_vars(self, x, force=False):
        x = _f64(x).ravel()
        ch = C.c_int(0)
        _chk(lib().orc_top_set_vars(self.h, _p(x), int(force), C.byref(ch)))
        return bool(ch.value)

    def top_objective(self):
        J = C.c_double(0)
        _chk(lib().orc_top_objective(self.h, C.byref(J)))
        return J.value

    def top_gradient(self):
        out = np.empty(self.ne)
        _chk(lib().orc_top_gradient(self.h, _p(out)))
        return out

    def top_constraint(self):
        c = C.c_double(0)
        _chk(lib().orc_top_constraint(self.h, C.byref(c)))
        return c.value

    def top_constraint_jac(self):
        out = np.empty(self.ne)
        _chk(lib().orc_top_constraint_jac(self.h, _p(out)))
        return out

    def top_u(self):
        out = np.empty((self.nn, self.dim))
        _chk(lib().orc_top_get_u(self.h, _p(out)))
        return out

    def top_vars(self, which=0):
        out = np.empty(self.ne)
        _chk(lib().orc_top_get_vars(self.h, int(which), _p(out)))
        return out

    def top_last_iters(self):
        return lib().orc_top_last_iters(self.h)

    def oc_step(self, m=0.2, ctol=1e-6):
        lam = C.c_double(0)
        nb = C.c_int(0)
        _chk(lib().orc_oc_step(self.h, C.c_double(m), C.c_double(ctol), C.byref(lam), C.byref(nb)))
        return lam.value, nb.value


def apply_filter(kind, param, dims, x, vars=None, backprop=False):
    dims = np.ascontiguousarray(dims, np.uint64)
    x = _f64(x).ravel()
    v = _f64(vars).ravel() if vars is not None else x
    out = np.empty_like(x)
    _chk(lib().orc_filter(0 if kind == "smoothing" else 1, C.c_double(param), _p(dims), len(dims), _p(x), _p(v), int(backprop), _p(out)))
    return out


def gauss_monomial(rule_degrees, exponents):
    rd = np.ascontiguousarray(rule_degrees, np.int32)
    ex = np.ascontiguousarray(exponents, np.int32)
    out = C.c_double(0)
    _chk(lib().orc_gauss_monomial(len(rd), _p(rd), _p(ex), C.byref(out)))
    return out.value


def set_num_threads(n):
    lib().orc_set_num_threads(int(n))


def max_threads():
    return lib().orc_max_threads()
