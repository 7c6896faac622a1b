"""Import shim for MeshFEM's `mesh` python module, restricted to what the ndr scripts use on the VoxelFEM path:
`mesh.MSHFieldWriter(path, V, F).addField(name, field)` (utils.py:315-326, checkpoints of the 3-D runs) and
`mesh.MSHFieldParser3(mshPath).scalarField(name)` (utils.py:410-415).

The file format is Gmsh MSH 2.2 exactly as MeshFEM writes it (MeshIO.cc:533-612 for the mesh, MSHFieldWriter.hh:128-196
for `$ElementData` / `$NodeData` sections; binary by default), so files written here load in the reference and vice versa.
Host-side I/O only: nothing here touches the device path."""
import enum
import struct

import numpy as np

# Gmsh element type <-> nodes per element (MeshIO.cc elementInfoForNodeCount)
# MESH_GUESS quirk kept for file compatibility: a 4-node element is written as a tet (type 4), also for the 2-D quad meshes
# of VoxelFEM ("we assume an element with 4 nodes is a tet, not a quad", MeshIO.cc:525-531)
_TYPE_FOR_NODES = {2: 1, 3: 2, 4: 4, 8: 5, 6: 9, 10: 11}
_NODES_FOR_TYPE = {1: 2, 2: 3, 3: 4, 4: 4, 5: 8, 9: 6, 11: 10}


class DomainType(enum.IntEnum):
    PER_ELEMENT = 0
    PER_NODE = 1
    GUESS = 2
    ANY = 3
    UNKNOWN = 4


class MSHFieldWriter:
    DomainType = DomainType

    def __init__(self, path, V, F, binary=True):
        V = np.asarray(V, dtype=np.float64)
        F = np.asarray(F, dtype=np.int64)
        if V.ndim != 2 or F.ndim != 2:
            raise RuntimeError("V must be (numVertices, 2|3) and F (numElements, nodesPerElement)")
        if V.shape[0] == 0:
            raise RuntimeError("Empty mesh.")
        if V.shape[1] == 2:
            V = np.hstack([V, np.zeros((V.shape[0], 1))])
        self._nv, self._ne, self._binary = V.shape[0], F.shape[0], bool(binary)
        if F.shape[1] not in _TYPE_FOR_NODES:
            raise RuntimeError("Illegal sized element (%d)" % F.shape[1])
        etype = _TYPE_FOR_NODES[F.shape[1]]
        try:
            self._f = open(path, "wb")
        except OSError:
            raise RuntimeError("Failed to open output file '%s'" % path)
        w = self._f.write
        w(b"$MeshFormat\n2.2 %d 8\n" % (1 if binary else 0))
        if binary:
            w(struct.pack("<i", 1) + b"\n")
        w(b"$EndMeshFormat\n$Nodes\n%d\n" % self._nv)
        if binary:
            rec = np.empty(self._nv, dtype=[("i", "<i4"), ("p", "<f8", (3,))])
            rec["i"] = np.arange(1, self._nv + 1)
            rec["p"] = V
            w(rec.tobytes() + b"\n")
        else:
            for i, p in enumerate(V):
                w(("%d %.17g %.17g %.17g\n" % (i + 1, p[0], p[1], p[2])).encode())
        w(b"$EndNodes\n$Elements\n%d\n" % self._ne)
        if binary:
            if self._ne:
                w(struct.pack("<iii", etype, self._ne, 0))
            rec = np.empty((self._ne, F.shape[1] + 1), dtype="<i4")
            rec[:, 0] = np.arange(1, self._ne + 1)
            rec[:, 1:] = F + 1
            w(rec.tobytes() + b"\n")
        else:
            for i, e in enumerate(F):
                w(("%d %d 0 %s\n" % (i + 1, etype, " ".join(str(int(v) + 1) for v in e))).encode())
        w(b"$EndElements\n")
        self._f.flush()

    def addField(self, name, field, dtype=DomainType.GUESS):
        f = np.asarray(field, dtype=np.float64)
        if f.ndim == 1:
            f = f[:, None]
        n, dim = f.shape
        dtype = DomainType(dtype)
        if dtype in (DomainType.GUESS, DomainType.ANY):  # MSHFieldWriter.hh:329-352: elements win when the counts tie
            dtype = DomainType.PER_ELEMENT if n == self._ne else DomainType.PER_NODE if n == self._nv else DomainType.UNKNOWN
        if (dtype == DomainType.PER_ELEMENT and n != self._ne) or (dtype == DomainType.PER_NODE and n != self._nv) or dtype == DomainType.UNKNOWN:
            raise RuntimeError("Invalid field domain size.")
        if dim == 2:  # 2-vectors are padded to 3-vectors for Gmsh
            f = np.hstack([f, np.zeros((n, 1))])
        elif dim == 4:  # flattened symmetric 2x2? not used on this path
            raise RuntimeError("Invalid field dimension.")
        if f.shape[1] not in (1, 3, 9):
            raise RuntimeError("Invalid field dimension.")
        sec = b"ElementData" if dtype == DomainType.PER_ELEMENT else b"NodeData"
        w = self._f.write
        w(b"$" + sec + b"\n1\n\"" + name.encode() + b"\"\n0\n3\n0\n%d\n%d\n" % (f.shape[1], n))
        if self._binary:
            rec = np.empty(n, dtype=[("i", "<i4"), ("v", "<f8", (f.shape[1],))])
            rec["i"] = np.arange(1, n + 1)
            rec["v"] = f
            w(rec.tobytes())
        else:
            for i, v in enumerate(f):
                w(("%d %s\n" % (i + 1, " ".join(repr(float(x)) for x in v))).encode())
        w(b"$End" + sec + b"\n")
        self._f.flush()

    def numVertices(self):
        return self._nv

    def numElements(self):
        return self._ne

    def close(self):
        if self._f:
            self._f.close()
            self._f = None

    def __del__(self):
        if getattr(self, "_f", None):
            self._f.close()


class _MSHFieldParser:
    """MSHFieldParser<N> (MSHFieldParser.cc): mesh + the named data sections of an MSH 2.2 file (ASCII or binary)."""
    _N = 3

    def __init__(self, mshPath):
        try:
            data = open(mshPath, "rb").read()
        except OSError:
            raise RuntimeError("Couldn't open input file " + str(mshPath))
        self._fields = {}
        pos = 0

        def line():
            nonlocal pos
            e = data.index(b"\n", pos)
            s = data[pos:e]
            pos = e + 1
            return s.strip()

        def expect(tag):
            got = line()
            while got == b"":
                got = line()
            if got != tag:
                raise RuntimeError("MSH parse error: expected %s, got %s" % (tag.decode(), got[:40].decode(errors="replace")))

        expect(b"$MeshFormat")
        ver, ftype, dsize = line().split()
        binary = int(ftype) == 1
        if binary:
            pos += 4 + 1  # the int 1 + newline
        expect(b"$EndMeshFormat")
        expect(b"$Nodes")
        nv = int(line())
        if binary:
            rec = np.frombuffer(data, dtype=[("i", "<i4"), ("p", "<f8", (3,))], count=nv, offset=pos)
            pos += rec.nbytes
            self._V = np.array(rec["p"])
        else:
            self._V = np.array([[float(t) for t in line().split()[1:4]] for _ in range(nv)]).reshape(nv, 3)
        expect(b"$EndNodes")
        expect(b"$Elements")
        ne = int(line())
        if binary:
            elems = []
            while len(elems) < ne:
                etype, cnt, ntags = struct.unpack_from("<iii", data, pos)
                pos += 12
                npe = _NODES_FOR_TYPE[etype]
                rec = np.frombuffer(data, dtype="<i4", count=cnt * (1 + ntags + npe), offset=pos).reshape(cnt, 1 + ntags + npe)
                pos += rec.nbytes
                elems.extend(rec[:, 1 + ntags:] - 1)
            self._F = np.array(elems, dtype=np.int64).reshape(ne, -1)
        else:
            rows = []
            for _ in range(ne):
                t = [int(v) for v in line().split()]
                rows.append([v - 1 for v in t[3 + t[2]:]])
            self._F = np.array(rows, dtype=np.int64).reshape(ne, -1)
        expect(b"$EndElements")
        while pos < len(data) and b"\n" in data[pos:]:
            hdr = line()
            if not hdr:
                continue
            if hdr not in (b"$ElementData", b"$NodeData"):
                end = b"$End" + hdr[1:]  # skip sections this shim does not model ($ElementNodeData, ...)
                pos = data.index(end, pos) + len(end)
                continue
            nstr = int(line())
            names = [line().strip(b'"').decode() for _ in range(nstr)]
            for _ in range(int(line())):
                line()
            itags = [int(line()) for _ in range(int(line()))]
            dim, n = itags[1], itags[2]
            if binary:
                rec = np.frombuffer(data, dtype=[("i", "<i4"), ("v", "<f8", (dim,))], count=n, offset=pos)
                pos += rec.nbytes
                vals = np.array(rec["v"]).reshape(n, dim)
            else:
                vals = np.array([[float(t) for t in line().split()[1:]] for _ in range(n)]).reshape(n, dim)
            expect(b"$End" + hdr[1:])
            self._fields[names[0]] = (DomainType.PER_ELEMENT if hdr == b"$ElementData" else DomainType.PER_NODE, vals)

    def vertices(self):
        return self._V[:, :self._N].copy() if self._N == 2 else self._V.copy()

    def elements(self):
        return self._F.copy()

    def _get(self, name, dim, domain):
        if name not in self._fields:
            raise RuntimeError("Couldn't find field " + name)
        dt, v = self._fields[name]
        if domain not in (DomainType.GUESS, DomainType.ANY) and DomainType(domain) != dt:
            raise RuntimeError("Couldn't find field " + name)
        if dim == 1 and v.shape[1] != 1:
            raise RuntimeError("Field %s is not a scalar field" % name)
        return v

    def scalarField(self, name, domain=DomainType.ANY):
        return self._get(name, 1, domain)[:, 0].copy()

    def vectorField(self, name, domain=DomainType.ANY):
        return self._get(name, 3, domain)[:, :self._N].copy()

    def scalarFieldNames(self):
        return [k for k, (_, v) in self._fields.items() if v.shape[1] == 1]

    def vectorFieldNames(self):
        return [k for k, (_, v) in self._fields.items() if v.shape[1] == 3]


class MSHFieldParser3(_MSHFieldParser):
    _N = 3


class MSHFieldParser2(_MSHFieldParser):
    _N = 2
