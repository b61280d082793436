"""MPAS-convention file I/O both ways, without a netCDF library: classic netCDF
files in the CDF-1 / CDF-2 (64-bit offset) / CDF-5 (64-bit data) variants.

Why here: the reference reads its mesh, initial state and restart through
SCORPIO / netCDF (src/base/IO.cpp, src/infra/IOStream.cpp), neither of which
exists on this box, and the state of BASELINE config 4 has variables beyond the
4 GiB limit of CDF-2 (2,621,442 cells x 80 levels x 8 B = 1.7 GB per field is
fine, but [tracer][cell][level] blocks and the 7.9 M-edge velocity are not once
tracers are stored as one variable) -- CDF-5 lifts that limit.  The writer lays the
file out up front and hands out big-endian ``numpy.memmap`` views, so a field can
be written slab by slab straight from ``omega_field_to_host`` output, and several
ranks can scatter their OWNED rows into one file by global ID.

Format (netCDF classic format specification; CDF-5 = PnetCDF's extension):
  header  = magic numrecs dim_list gatt_list var_list
  magic   = 'C' 'D' 'F' version(1|2|5)
  NON_NEG = INT32 (CDF-1/2) | INT64 (CDF-5); OFFSET = INT32 (CDF-1) | INT64 (CDF-2/5)
  dim     = name NON_NEG(length);  var = name NON_NEG(ndims) dimid* vatt_list nc_type NON_NEG(vsize) OFFSET(begin)
  data is big-endian, every fixed-size variable padded to 4 bytes.
Record (unlimited) dimensions are not used by Omega's streams and not supported.

Restart contents follow src/ocn/OceanInit.cpp:152-177 and configs/Default.yml:76-131:
the State group (NormalVelocity, LayerThickness), the tracer arrays by name and
the SimulationTime metadata as a global attribute.
"""
from __future__ import annotations

import os
import struct

import numpy as np

NC_BYTE, NC_CHAR, NC_SHORT, NC_INT, NC_FLOAT, NC_DOUBLE = 1, 2, 3, 4, 5, 6
NC_DIMENSION, NC_VARIABLE, NC_ATTRIBUTE = 0x0A, 0x0B, 0x0C
_DT = {NC_BYTE: ">i1", NC_CHAR: "S1", NC_SHORT: ">i2", NC_INT: ">i4", NC_FLOAT: ">f4", NC_DOUBLE: ">f8"}
_TYPE_OF = {"i4": NC_INT, "f8": NC_DOUBLE, "f4": NC_FLOAT, "i2": NC_SHORT, "i1": NC_BYTE}


def _pad4(n):
    return (n + 3) & ~3


class CDFWriter:
    """Define dimensions / attributes / variables, then ``create()`` and fill ``var(name)[...]``."""

    def __init__(self, path: str, version: int = 5):
        assert version in (1, 2, 5)
        self.path, self.version = path, version
        self.dims, self.atts, self.vars = [], [], []   # (name, len) | (name, value) | dict
        self._layout = None

    # -- definition ---------------------------------------------------------------
    def dim(self, name, length):
        self.dims.append((name, int(length)))

    def att(self, name, value):
        self.atts.append((name, value))

    def variable(self, name, dtype, dims, atts=None):
        self.vars.append(dict(name=name, type=_TYPE_OF[dtype], dims=list(dims), atts=list((atts or {}).items())))

    # -- encoding -------------------------------------------------------------------
    def _nn(self, v):
        return struct.pack(">q" if self.version == 5 else ">i", v)

    def _off(self, v):
        return struct.pack(">i" if self.version == 1 else ">q", v)

    def _name(self, s):
        b = s.encode()
        return self._nn(len(b)) + b + b"\0" * (_pad4(len(b)) - len(b))

    def _att_list(self, atts):
        if not atts:
            return struct.pack(">i", 0) + self._nn(0)
        out = struct.pack(">i", NC_ATTRIBUTE) + self._nn(len(atts))
        for name, v in atts:
            out += self._name(name)
            if isinstance(v, (bytes, str)):
                b = v.encode() if isinstance(v, str) else v
                out += struct.pack(">i", NC_CHAR) + self._nn(len(b)) + b + b"\0" * (_pad4(len(b)) - len(b))
            else:
                a = np.atleast_1d(np.asarray(v))
                t = NC_INT if a.dtype.kind in "iu" else NC_DOUBLE
                raw = a.astype(_DT[t]).tobytes()
                out += struct.pack(">i", t) + self._nn(a.size) + raw + b"\0" * (_pad4(len(raw)) - len(raw))
        return out

    def _header(self, begins):
        dimid = {n: i for i, (n, _) in enumerate(self.dims)}
        h = b"CDF" + bytes([self.version]) + self._nn(0)
        h += struct.pack(">i", NC_DIMENSION) + self._nn(len(self.dims)) if self.dims else struct.pack(">i", 0) + self._nn(0)
        for n, l in self.dims:
            h += self._name(n) + self._nn(l)
        h += self._att_list(self.atts)
        h += struct.pack(">i", NC_VARIABLE) + self._nn(len(self.vars)) if self.vars else struct.pack(">i", 0) + self._nn(0)
        for v, b in zip(self.vars, begins):
            h += self._name(v["name"]) + self._nn(len(v["dims"]))
            for d in v["dims"]:
                h += self._nn(dimid[d])
            h += self._att_list(v["atts"])
            vsize = v["vsize"]
            if self.version != 5 and vsize > 0xFFFFFFFF - 3:
                vsize = 0xFFFFFFFF   # classic formats: "vsize is redundant" once it does not fit
            h += struct.pack(">i", v["type"]) + (self._nn(vsize) if self.version == 5 else struct.pack(">I", vsize)) + self._off(b)
        return h

    def create(self):
        dlen = dict(self.dims)
        for v in self.vars:
            v["shape"] = tuple(dlen[d] for d in v["dims"])
            n = int(np.prod(v["shape"], dtype=np.int64)) if v["shape"] else 1
            v["nbytes"] = n * np.dtype(_DT[v["type"]]).itemsize
            v["vsize"] = _pad4(v["nbytes"])
        if self.version != 5:
            big = [v["name"] for v in self.vars[:-1] if v["vsize"] > 0xFFFFFFFF - 3]
            if big or any(l > 0x7FFFFFFF for _, l in self.dims):
                raise ValueError(f"variables {big} exceed the 4 GiB limit of CDF-{self.version}: use version=5")
        hlen = len(self._header([0] * len(self.vars)))
        begins, pos = [], _pad4(hlen)
        for v in self.vars:
            begins.append(pos)
            pos += v["vsize"]
        if self.version == 1 and pos > 0x7FFFFFFF:
            raise ValueError("file exceeds the 2 GiB limit of CDF-1")
        header = self._header(begins)
        assert len(header) == hlen
        with open(self.path, "wb") as f:
            f.write(header)
            f.truncate(max(pos, 1))
        self._layout = {v["name"]: (b, v) for v, b in zip(self.vars, begins)}
        return self

    def var(self, name, mode="r+"):
        """big-endian memmap view of a variable (assign native arrays into it)."""
        b, v = self._layout[name]
        return np.memmap(self.path, dtype=_DT[v["type"]], mode=mode, offset=b, shape=v["shape"] or (1,))


class CDFReader:
    """Header of a CDF-1/2/5 file; ``var(name)`` gives a read-only big-endian memmap."""

    def __init__(self, path: str):
        self.path = path
        with open(path, "rb") as f:
            head = f.read(4)
            if head[:3] != b"CDF" or head[3] not in (1, 2, 5):
                raise ValueError(f"{path}: not a classic netCDF file (magic {head!r})")
            self.version = head[3]
            # headers are small; read generously and parse from memory
            buf = head + f.read(1 << 22)
        self._b, self._p = buf, 4
        nn = self._nn
        nn()  # numrecs
        self.dims, self.atts, self.vars = [], {}, {}
        tag = self._i4()
        n = nn()
        assert tag in (0, NC_DIMENSION)
        for _ in range(n):
            name = self._name()
            self.dims.append((name, nn()))
        self.atts = self._att_list()
        tag = self._i4()
        n = nn()
        assert tag in (0, NC_VARIABLE)
        for _ in range(n):
            name = self._name()
            nd = nn()
            dimids = [nn() for _ in range(nd)]
            atts = self._att_list()
            t = self._i4()
            nn() if self.version == 5 else self._i4()           # vsize (redundant)
            begin = struct.unpack_from(">i" if self.version == 1 else ">q", self._b, self._p)[0]
            self._p += 4 if self.version == 1 else 8
            shape = tuple(self.dims[d][1] for d in dimids)
            self.vars[name] = dict(type=t, dims=[self.dims[d][0] for d in dimids], shape=shape, begin=begin, atts=atts)
        del self._b

    def _i4(self):
        v = struct.unpack_from(">i", self._b, self._p)[0]
        self._p += 4
        return v

    def _nn(self):
        if self.version == 5:
            v = struct.unpack_from(">q", self._b, self._p)[0]
            self._p += 8
        else:
            v = struct.unpack_from(">i", self._b, self._p)[0]
            self._p += 4
        return v

    def _name(self):
        n = self._nn()
        s = self._b[self._p:self._p + n].decode()
        self._p += _pad4(n)
        return s

    def _att_list(self):
        tag = self._i4()
        n = self._nn()
        assert tag in (0, NC_ATTRIBUTE)
        out = {}
        for _ in range(n):
            name = self._name()
            t = self._i4()
            cnt = self._nn()
            size = cnt * np.dtype(_DT[t]).itemsize
            raw = self._b[self._p:self._p + size]
            self._p += _pad4(size)
            out[name] = raw.decode() if t == NC_CHAR else np.frombuffer(raw, _DT[t]).astype(_DT[t][1:])
        return out

    def dim(self, name):
        return dict(self.dims)[name]

    def var(self, name):
        v = self.vars[name]
        return np.memmap(self.path, dtype=_DT[v["type"]], mode="r", offset=v["begin"], shape=v["shape"] or (1,))

    def read(self, name):
        a = self.var(name)
        return np.array(a).astype(a.dtype.newbyteorder("="))


# ---- restart files (State + tracers + SimulationTime) -------------------------------------
TRACER_NAMES = ("Temperature", "Salinity", "Debug1", "Debug2", "Debug3")   # src/ocn/TracerDefs.inc:24-41


def tracer_name(l):
    return TRACER_NAMES[l] if l < len(TRACER_NAMES) else f"Tracer{l}"


def sim_time_string(seconds: float, year0: int = 1) -> str:
    """'YYYY-MM-DD_HH:MM:SS' in the No-Leap calendar (configs/Default.yml:8), elapsed since 0001-01-01."""
    s = int(round(seconds))
    day, s = divmod(s, 86400)
    year, doy = divmod(day, 365)
    mdays = (31, 28, 31, 30, 31, 30, 31, 31, 30, 31, 30, 31)
    mon = 0
    while doy >= mdays[mon]:
        doy -= mdays[mon]
        mon += 1
    return f"{year0 + year:04d}-{mon + 1:02d}-{doy + 1:02d}_{s // 3600:02d}:{s % 3600 // 60:02d}:{s % 60:02d}"


def create_restart(path, n_cells, n_edges, n_levels, n_tracers, sim_seconds, version=5):
    """Lay out a restart file for the GLOBAL mesh (one rank calls this, every rank then scatters its owned rows)."""
    w = CDFWriter(path, version)
    w.dim("NCells", n_cells)
    w.dim("NEdges", n_edges)
    w.dim("NVertLayers", n_levels)
    w.att("SimulationTime", sim_time_string(sim_seconds))
    w.att("SimulationSeconds", float(sim_seconds))
    w.att("file_id", "omega_b200 restart")
    w.variable("NormalVelocity", "f8", ("NEdges", "NVertLayers"), {"units": "m s-1"})
    w.variable("LayerThickness", "f8", ("NCells", "NVertLayers"), {"units": "m"})
    for l in range(n_tracers):
        w.variable(tracer_name(l), "f8", ("NCells", "NVertLayers"))
    return w.create()


def scatter_owned(path, L, h, u, tr):
    """Write one rank's OWNED rows (host order arrays of the local mesh L) at their global IDs."""
    r = CDFReader(path)
    nco, neo = int(L.NCellsOwned), int(L.NEdgesOwned)
    cid, eid = np.asarray(L.CellID[:nco]) - 1, np.asarray(L.EdgeID[:neo]) - 1

    def put(name, ids, rows):
        v = r.vars[name]
        mm = np.memmap(path, dtype=_DT[v["type"]], mode="r+", offset=v["begin"], shape=v["shape"])
        order = np.argsort(ids, kind="stable")       # ascending file offsets
        mm[ids[order]] = rows[order]
        mm.flush()

    put("LayerThickness", cid, h[:nco])
    put("NormalVelocity", eid, u[:neo])
    for l in range(tr.shape[0]):
        put(tracer_name(l), cid, tr[l, :nco])


def read_restart(path, L, n_tracers):
    """(h, u, tracers, SimulationTime, seconds) for the local mesh L in host order, halo rows included
    (the reference fills them with an exchange after the read, OceanInit.cpp:179-186; reading them by global
    ID is the same data)."""
    r = CDFReader(path)
    nca, nea = int(L.NCellsAll), int(L.NEdgesAll)
    K = r.dim("NVertLayers")
    cid, eid = np.asarray(L.CellID[:nca]) - 1, np.asarray(L.EdgeID[:nea]) - 1

    def get(name, ids, n_all):
        out = np.zeros((n_all + 1, K))
        out[:n_all] = np.asarray(r.var(name)[ids])
        return out

    h = get("LayerThickness", cid, nca)
    u = get("NormalVelocity", eid, nea)
    tr = np.zeros((n_tracers, nca + 1, K))
    for l in range(n_tracers):
        tr[l] = get(tracer_name(l), cid, nca)
    secs = float(np.atleast_1d(r.atts.get("SimulationSeconds", np.array([0.0])))[0])
    return h, u, tr, r.atts.get("SimulationTime", " "), secs
