"""Python mirror of the reference's `pyVoxelFEM` pybind11 module (VoxelFEM/python_bindings/VoxelFEM.cc)
over the C ABI of libvfem_b200.so.  Same class / function / attribute names and argument meaning, so
fem.py, helpers/ipopt_helpers.py and the training scripts run unchanged against it.

numpy float64 arrays are copied through (like pybind's Eigen casters); CUDA tensors (anything with
`__cuda_array_interface__` or `__dlpack__`, e.g. torch tensors) are used in place - the zero-copy path
for the neural-field loop (`setVars` / `evaluateObjectiveGradient(out=...)`).
"""
import ctypes as C
import json

import numpy as np

from ._lib import FILTER_CB, PCG_CB, RESIDUAL_CB, SOLVE_CB, check, lib
from .benchmark import timed as _timed

HOST, DEVICE = 0, 1


def _np(a, dtype=np.float64):
    return np.ascontiguousarray(a, dtype=dtype)


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _is_device(a):
    return hasattr(a, "__cuda_array_interface__") and not isinstance(a, np.ndarray)


def _dev_ptr(a, count):
    """Raw device pointer of a contiguous float64 CUDA array with `count` entries."""
    cai = a.__cuda_array_interface__
    if cai["typestr"] not in ("<f8", "=f8", "f8"):
        raise TypeError("device arrays must be float64 (got %s)" % cai["typestr"])
    if cai.get("strides") is not None:
        raise ValueError("device arrays must be contiguous")
    n = 1
    for s in cai["shape"]:
        n *= s
    if n != count:
        raise ValueError("device array has %d entries, expected %d" % (n, count))
    return C.c_void_p(cai["data"][0])


class _Arg:
    """Resolves an input array to (pointer, memspace); keeps host staging alive."""

    def __init__(self, a, count):
        if _is_device(a):
            self.ms, self.keep, self.ptr = DEVICE, a, _dev_ptr(a, count)
        else:
            arr = _np(a).reshape(-1)
            if arr.size != count:
                raise RuntimeError("array has %d entries, expected %d" % (arr.size, count))
            self.ms, self.keep, self.ptr = HOST, arr, _ptr(arr)


def parse_bc_regions(path, bbmin, bbmax, dim):
    """Subset of MeshFEM's .bc JSON format used by the shipped files (BoundaryConditions.cc:218-388):
    `box` / `box%` regions, types `force` and `dirichlet[xyz]*`, numeric values."""
    with open(path) as f:
        params = json.load(f)
    bbmin, bbmax = np.asarray(bbmin, float), np.asarray(bbmax, float)
    out = []
    for r in params["regions"]:
        typ = r["type"]
        comp = np.ones(dim, np.uint8)
        if typ.startswith("dirichlet"):
            letters = ""
            for ch in typ[len("dirichlet"):]:
                if ch not in "xyz":
                    break
                letters += ch
            if len(letters) > 3:
                raise RuntimeError("Invalid dirichlet component mask")
            if letters:
                comp = np.zeros(dim, np.uint8)
                for ch in letters:
                    if "xyz".index(ch) < dim:
                        comp[("xyz").index(ch)] = 1
            typ = "dirichlet" + typ[len("dirichlet") + len(letters):]
        if "box" in r:
            lo, hi = (np.array(r["box"][k], float)[:dim] for k in ("minCorner", "maxCorner"))
        elif "box%" in r:  # relative to the domain bounding box (Geometry.hh:259-262)
            lo, hi = (bbmin + np.array(r["box%"][k], float)[:dim] * (bbmax - bbmin) for k in ("minCorner", "maxCorner"))
        else:
            raise RuntimeError("Unsupported region geometry (only box / box% regions are implemented)")
        val = np.array(r["value"], float)[:dim]
        if typ not in ("force", "dirichlet"):
            raise RuntimeError("Invalid boundary condition type: " + r["type"])
        out.append((typ, _np(lo), _np(hi), comp, _np(val)))
    return out


class TensorProductSimulator:
    """pyVoxelFEM.TensorProductSimulator(degreesPerDimension, domainBBox, elementsPerDimension) (VoxelFEM.cc:234-240)."""

    def __init__(self, degreesPerDimension, domainBBox, elementsPerDimension):
        deg = _np(degreesPerDimension, np.uint64)
        el = _np(elementsPerDimension, np.uint64)
        self._bbmin, self._bbmax = _np(domainBBox[0]), _np(domainBBox[1])
        self.N = len(el)
        if len(deg) != self.N:
            raise RuntimeError("No template instantiation matching degreesPerDimension!")
        self._h = C.c_void_p()
        check(lib().vfem_tps_create(self.N, _ptr(deg), _ptr(self._bbmin), _ptr(self._bbmax), _ptr(el), C.byref(self._h)))
        self._elems = [int(e) for e in el]

    def __del__(self):
        if getattr(self, "_h", None) and lib is not None:  # `lib` is None during interpreter shutdown
            try:
                lib().vfem_tps_destroy(self._h)
            except Exception:  # interpreter shutdown
                pass
            self._h = None

    # --- sizes / indexing (NDVector order: first axis slowest) ---
    def numNodes(self):
        return lib().vfem_tps_num_nodes(self._h)

    def numElements(self):
        return lib().vfem_tps_num_elements(self._h)

    def numDoFs(self):
        return self.numNodes() * self.N

    def NbElementsPerDimension(self):
        return list(self._elems)

    def nodePosition(self, ni):
        p = np.empty(self.N)
        check(lib().vfem_tps_node_position(self._h, int(ni), _ptr(p)))
        return p

    def elementIndexForGridCell(self, cellIdxs):
        ei = 0
        for d, c in enumerate(cellIdxs):
            ei = ei * self._elems[d] + int(c)
        return ei

    def elemNodeGlobalIndex(self, ei, n):
        idx = np.unravel_index(int(ei), self._elems)
        nper = [e + 1 for e in self._elems]
        loc = [(int(n) >> (self.N - 1 - d)) & 1 for d in range(self.N)]
        return int(np.ravel_multi_index([i + l for i, l in zip(idx, loc)], nper))

    def elementNodes(self, ei):
        return [self.elemNodeGlobalIndex(ei, n) for n in range(2 ** self.N)]

    # --- material / SIMP ---
    def readMaterial(self, materialPath):
        with open(materialPath) as f:
            m = json.load(f)
        if m.get("type") != "isotropic_material":
            raise RuntimeError("only isotropic_material files are supported")
        check(lib().vfem_tps_set_isotropic(self._h, float(m["young"]), float(m["poisson"])))

    def setIsotropicMaterial(self, young, poisson):
        check(lib().vfem_tps_set_isotropic(self._h, float(young), float(poisson)))

    def _simp(self):
        a, b, c = C.c_double(), C.c_double(), C.c_double()
        check(lib().vfem_tps_get_simp(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return [a.value, b.value, c.value]

    def _set_simp(self, i, v):
        s = self._simp()
        s[i] = float(v)
        check(lib().vfem_tps_set_simp(self._h, *s))

    E_0 = property(lambda self: self._simp()[0], lambda self, v: self._set_simp(0, v))
    E_min = property(lambda self: self._simp()[1], lambda self, v: self._set_simp(1, v))
    gamma = property(lambda self: self._simp()[2], lambda self, v: self._set_simp(2, v))

    def fullDensityElementStiffnessMatrix(self):
        ks = self.N * 2 ** self.N
        k0 = np.empty((ks, ks))
        check(lib().vfem_tps_get_k0(self._h, _ptr(k0)))
        return k0

    def elementStiffnessMatrix(self, ei):
        rho = self.elementDensity(ei)
        E0, Emin, g = self._simp()
        return (Emin + rho ** g * (E0 - Emin)) * self.fullDensityElementStiffnessMatrix()

    def clearCachedElementStiffness(self):
        pass  # the device path never caches per-element matrices on the fine level

    # --- boundary conditions ---
    def applyDisplacementsAndLoadsFromFile(self, bcPath):
        for typ, lo, hi, comp, val in parse_bc_regions(bcPath, self._bbmin, self._bbmax, self.N):
            if typ == "force":
                check(lib().vfem_tps_add_force_region(self._h, _ptr(lo), _ptr(hi), _ptr(val)))
            else:
                check(lib().vfem_tps_add_dirichlet_region(self._h, _ptr(lo), _ptr(hi), _ptr(comp), _ptr(val)))

    def _get_dirichlet(self):
        m = np.empty((self.numNodes(), self.N), np.uint8)
        v = np.empty((self.numNodes(), self.N))
        check(lib().vfem_tps_get_dirichlet(self._h, _ptr(m), _ptr(v)))
        return m.astype(bool), v

    def _set_dirichlet_mask(self, m):
        check(lib().vfem_tps_set_dirichlet(self._h, _ptr(_np(m, np.uint8)), None))

    def _set_dirichlet_values(self, v):
        check(lib().vfem_tps_set_dirichlet(self._h, None, _ptr(_np(v))))

    dirichletMask = property(lambda self: self._get_dirichlet()[0], _set_dirichlet_mask)
    dirichletValues = property(lambda self: self._get_dirichlet()[1], _set_dirichlet_values)

    def getDirichletVarsAndValues(self):
        m, v = self._get_dirichlet()
        idx = np.flatnonzero(m.reshape(-1))
        return list(idx), list(v.reshape(-1)[idx])

    def getForceMask(self):
        m = np.empty(self.numNodes(), np.uint8)
        check(lib().vfem_tps_get_force_mask(self._h, _ptr(m)))
        return m.astype(bool)

    def buildLoadVector(self):
        f = np.empty((self.numNodes(), self.N))
        check(lib().vfem_tps_build_load_vector(self._h, _ptr(f), HOST))
        return f

    # --- densities ---
    def getDensities(self):
        rho = np.empty(self.numElements())
        check(lib().vfem_tps_get_densities(self._h, _ptr(rho), HOST))
        return rho

    def setElementDensities(self, rho):
        a = _Arg(rho, self.numElements())
        check(lib().vfem_tps_set_densities(self._h, a.ptr, a.ms))

    def setUniformDensities(self, density):
        check(lib().vfem_tps_set_uniform_densities(self._h, float(density)))

    def setElementDensity(self, ei, value):
        check(lib().vfem_tps_set_element_density(self._h, int(ei), float(value)))

    def elementDensity(self, ei):
        v = C.c_double()
        check(lib().vfem_tps_element_density(self._h, int(ei), C.byref(v)))
        return v.value

    # --- operators ---
    @_timed(lambda self: "applyK " + "x".join(str(int(e)) for e in self._elems))
    def applyK(self, u, out=None):
        a = _Arg(u, self.numDoFs())
        if a.ms == DEVICE:
            if out is None:
                raise RuntimeError("device-resident applyK needs a preallocated `out` device array")
            check(lib().vfem_tps_apply_k(self._h, a.ptr, _dev_ptr(out, self.numDoFs()), DEVICE))
            return out
        res = np.empty((self.numNodes(), self.N)) if out is None else out
        if res.dtype != np.float64 or not res.flags.c_contiguous or res.size != self.numDoFs():
            raise RuntimeError("`out` must be a C-contiguous float64 array with numNodes*N entries")
        check(lib().vfem_tps_apply_k(self._h, a.ptr, _ptr(res), HOST))
        return res

    # kernel selection of the fine-level apply (not in the reference): 0 = auto, 1 = per-node gather kernel
    def setApplyVariant(self, variant):
        check(lib().vfem_tps_set_apply_variant(self._h, int(variant)))

    def applyInfo(self):
        info = (C.c_int * 7)()
        check(lib().vfem_tps_apply_info(self._h, info))
        keys = ("march", "k0_structure", "tile_shape", "layers_per_chunk", "grid_x", "grid_y", "grid_z")
        return dict(zip(keys, list(info)))

    def complianceGradient(self, u):
        a = _Arg(u, self.numDoFs())
        g = np.empty(self.numElements())
        if a.ms == DEVICE:
            raise RuntimeError("use ComplianceObjective.gradient(out=...) for device arrays")
        check(lib().vfem_tps_compliance_gradient(self._h, a.ptr, _ptr(g), HOST))
        return g

    def multigridSolver(self, numCoarseningLevels):
        from .voxelfem_mg import MultigridSolver
        return MultigridSolver(self, numCoarseningLevels)

    # --- mesh export / density import (host-side I/O: TPS.hh:470-492, 531-565) ---
    def _node_positions(self):
        nper = [e + 1 for e in self._elems]
        idx = np.indices(nper).reshape(self.N, -1).T
        spacing = (self._bbmax - self._bbmin) / np.array(self._elems, float)   # TPS.hh:274-279 (size / (n - 1))
        return self._bbmin + idx * spacing

    def getMesh(self):
        """(V, F): node positions (numNodes, 3) and element corner indices in Gmsh ordering (TPS.hh:531-565)."""
        P = self._node_positions()
        V = np.zeros((P.shape[0], 3))
        V[:, :self.N] = P
        nper = [e + 1 for e in self._elems]
        eidx = np.indices(self._elems).reshape(self.N, -1)
        npe = 2 ** self.N
        F = np.empty((eidx.shape[1], npe), np.int32)
        cols = []
        for n in range(npe):
            loc = [(n >> (self.N - 1 - d)) & 1 for d in range(self.N)]
            cols.append(np.ravel_multi_index([eidx[d] + loc[d] for d in range(self.N)], nper))
        k = 0
        for pair in range(npe // 2):  # counterclockwise on quads / on both quad faces of a hex
            a, b = cols[2 * pair], cols[2 * pair + 1]
            F[:, k], F[:, k + 1] = (a, b) if pair % 2 == 0 else (b, a)
            k += 2
        return V, F

    def readDensities(self, materialPath, fieldName="density"):
        """Per-element scalar field of a .msh file -> element densities, elements matched by their centre (TPS.hh:470-509)."""
        import os
        if os.path.splitext(materialPath)[1] != ".msh":
            raise RuntimeError("Material file extension" + os.path.splitext(materialPath)[1] + " is not supported")
        from .shims import mesh as _mesh
        fp = (_mesh.MSHFieldParser3 if self.N == 3 else _mesh.MSHFieldParser2)(materialPath)
        dens = fp.scalarField(fieldName, _mesh.DomainType.PER_ELEMENT)
        elems, verts = fp.elements(), fp._V
        if len(elems) != self.numElements():
            raise RuntimeError("The number of elements in the mesh : %d and the number of elements of the simulator : %d must be equal."
                               % (len(elems), self.numElements()))
        lo, hi = verts.min(axis=0), verts.max(axis=0)
        centre = verts[elems].mean(axis=1)
        rel = (centre - lo)[:, :self.N] / np.where(hi - lo > 0, hi - lo, 1.0)[:self.N]
        E = np.array(self._elems)
        cell = np.minimum(np.floor(rel * E).astype(np.int64), E - 1)
        rho = self.getDensities()
        rho[np.ravel_multi_index(cell.T, self._elems)] = dens
        self.setElementDensities(rho)

    def solve(self, f, out=None):
        """TPS::solve (TPS.hh:834-865): K(rho) u = f with the Dirichlet rows / columns removed; the factorisation is redone only
        when the densities, the material or the Dirichlet mask changed.  Non-zero Dirichlet values throw, as in the reference."""
        a = _Arg(f, self.numDoFs())
        if a.ms == DEVICE:
            if out is None:
                raise RuntimeError("device-resident solve needs a preallocated `out` device array")
            check(lib().vfem_tps_solve(self._h, a.ptr, _dev_ptr(out, self.numDoFs()), DEVICE))
            return out
        u = np.empty((self.numNodes(), self.N))
        check(lib().vfem_tps_solve(self._h, a.ptr, _ptr(u), HOST))
        return u

    def solveWithImposedLoads(self):
        return self.solve(self.buildLoadVector())   # TPS.hh:838-840

    def constantStrainLoad(self, eps):
        """Global load under the constant unit strain `eps` (TPS.hh:792-820 with Element::constantStrainLoad :167-173):
        per element rho_e * vol * int eps(phi_j e_i) : C : eps = rho_e * (K0 u_eps)|_e for the affine field u_eps(x) = eps x
        (scaled by the DENSITY, not the SIMP modulus, as in the reference).  Host-side helper of the homogenisation scripts."""
        eps = _np(eps).reshape(self.N, self.N)
        _, F = self._element_nodes()
        ue = (self._node_positions() @ eps.T)[F].reshape(F.shape[0], -1)
        le = (ue @ self.fullDensityElementStiffnessMatrix().T) * self.getDensities()[:, None]
        out = np.zeros((self.numNodes(), self.N))
        np.add.at(out, F.ravel(), le.reshape(-1, self.N))
        return out

    def getK(self):
        """Assembled stiffness matrix (TPS.hh:966-973 -> m_assembleStiffnessMatrix :589-625), Dirichlet rows/columns included, as a
        scipy.sparse CSC matrix (the reference returns MeshFEM's SuiteSparseMatrix; host-side export, not on the solve path)."""
        import scipy.sparse as sp
        K0 = self.fullDensityElementStiffnessMatrix()
        rho = self.getDensities()
        E0, Emin, g = self._simp()
        Ee = Emin + rho ** g * (E0 - Emin)
        _, F = self._element_nodes()
        dofs = (F[:, :, None] * self.N + np.arange(self.N)[None, None, :]).reshape(F.shape[0], -1)
        ks = dofs.shape[1]
        rows = np.repeat(dofs, ks, axis=1).ravel()
        cols = np.tile(dofs, (1, ks)).ravel()
        vals = (Ee[:, None] * K0.ravel()[None, :]).ravel()
        n = self.numDoFs()
        return sp.csc_matrix((vals, (rows, cols)), shape=(n, n))

    def _element_nodes(self):
        nper = [e + 1 for e in self._elems]
        eidx = np.indices(self._elems).reshape(self.N, -1)
        cols = []
        for n in range(2 ** self.N):
            loc = [(n >> (self.N - 1 - d)) & 1 for d in range(self.N)]
            cols.append(np.ravel_multi_index([eidx[d] + loc[d] for d in range(self.N)], nper))
        return eidx, np.stack(cols, axis=1)
