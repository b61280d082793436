"""Multi-GPU host logic: exchange lists and the halo exchange transport.

One process per GPU; the horizontal mesh is partitioned (decomp.py), each rank
steps its partition with the CUDA library and exchanges halos at the two
exchange points of an RK4 step (one for RK2 / forward-backward).

Exchange lists follow the reference (src/base/Halo.cpp:455-600): for each
index space the halo elements, visited layer by layer in ascending local
index, are appended to the receive list of their owner; the owner-local
indices (Loc(i,1)) form the owner's send list in the same order; neighbours
are the tasks that own my halo elements or need mine, ascending.  Buffer
order is [outer][exchange index][level] (src/base/Halo.h:304-502).

Transport: torch.distributed P2P (NCCL send/recv over NVLink; gloo on CPU for
the tests).  Pack / unpack are CUDA kernels of the library
(omega_halo_pack / omega_halo_unpack); the numpy versions below serve the
CPU tests and the in-process multi-rank emulation only.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi, decomp

KINDS = ("Cell", "Edge", "Vertex")


def _n(L, kind, what):
    return int(L["N%ss%s" % ("Vertice" if kind == "Vertex" else kind, what)])


def build_halo_lists(locs):
    """Per task: dict(neighbors=[...], send={kind: {nbr: idx}}, recv={kind: {nbr: idx}}) in host order."""
    nt = len(locs)
    out = [dict(neighbors=set(), send={k: {} for k in KINDS}, recv={k: {} for k in KINDS}) for _ in range(nt)]
    for t, L in enumerate(locs):
        for kind in KINDS:
            loc = L[kind + "Loc"]
            n_owned, n_all = _n(L, kind, "Owned"), _n(L, kind, "All")
            idx = np.arange(n_owned, n_all)
            owner = loc[idx, 0]
            for nbr in np.unique(owner):
                if nbr == t:
                    continue
                mine = idx[owner == nbr]                       # ascending local index
                out[t]["recv"][kind][int(nbr)] = mine.astype(np.int32)
                out[int(nbr)]["send"][kind][t] = loc[mine, 1].astype(np.int32)  # owner-local indices, same order
                out[t]["neighbors"].add(int(nbr))
                out[int(nbr)]["neighbors"].add(t)
    for o in out:
        o["neighbors"] = sorted(o["neighbors"])
    return out


def flatten_lists(h):
    """CSR form for OmegaHaloDesc: per kind (offsets[NN+1], idx) for send and recv."""
    nn = len(h["neighbors"])
    flat = {}
    for d in ("send", "recv"):
        for kind in KINDS:
            off = np.zeros(nn + 1, np.int32)
            parts = []
            for i, nbr in enumerate(h["neighbors"]):
                a = h[d][kind].get(nbr, np.zeros(0, np.int32))
                parts.append(a)
                off[i + 1] = off[i] + a.size
            flat[(d, kind)] = (off, np.concatenate(parts).astype(np.int32) if parts else np.zeros(0, np.int32))
    return flat


def local_problem(mesh, world, rank, nlev, halo_width=decomp.HALO_WIDTH):
    """(local mesh of `rank`, its halo lists).  Every rank derives the same
    partition deterministically, no communication needed."""
    locs = decomp.build_all(mesh, world, nlev, halo_width)
    lists = build_halo_lists(locs)
    return locs[rank], lists[rank]


# ---- numpy pack / unpack (tests, emulation) ----------------------------------
def pack_numpy(arr, idx):
    """arr [..., N, K] -> [outer][len(idx)][K] flattened (Halo.h:304-502 buffer order)."""
    a = arr.reshape((-1,) + arr.shape[-2:]) if arr.ndim >= 2 else arr.reshape(1, -1, 1)
    return np.ascontiguousarray(a[:, idx, :]).reshape(-1)


def unpack_numpy(arr, idx, buf):
    a = arr.reshape((-1,) + arr.shape[-2:]) if arr.ndim >= 2 else arr.reshape(1, -1, 1)
    a[:, idx, :] = buf.reshape(a.shape[0], idx.size, a.shape[2])


def exchange_inprocess(lists, kind_name, arrays):
    """arrays[t] = rank t's host-order array; copies owner rows into every halo row."""
    for t, h in enumerate(lists):
        for nbr in h["neighbors"]:
            ridx = h["recv"][kind_name].get(nbr)
            if ridx is None or ridx.size == 0:
                continue
            sidx = lists[nbr]["send"][kind_name][t]
            unpack_numpy(arrays[t], ridx, pack_numpy(arrays[nbr], sidx))


class HostHaloExchanger:
    """Halo::exchangeFullArrayHalo over torch.distributed on CPU tensors (gloo)."""

    def __init__(self, lists):
        self.h = lists

    def exchange(self, arr, kind_name):
        import torch
        import torch.distributed as dist
        ops, recvs = [], []
        for nbr in self.h["neighbors"]:
            sidx = self.h["send"][kind_name].get(nbr)
            ridx = self.h["recv"][kind_name].get(nbr)
            if sidx is not None and sidx.size:
                ops.append(dist.P2POp(dist.isend, torch.from_numpy(pack_numpy(arr, sidx)), nbr))
            if ridx is not None and ridx.size:
                outer = int(np.prod(arr.shape[:-2])) if arr.ndim > 2 else 1
                buf = torch.empty(outer * ridx.size * arr.shape[-1], dtype=torch.float64)
                recvs.append((ridx, buf))
                ops.append(dist.P2POp(dist.irecv, buf, nbr))
        if ops:
            for r in dist.batch_isend_irecv(ops):
                r.wait()
        for ridx, buf in recvs:
            unpack_numpy(arr, ridx, buf.numpy())
        return 0


class HaloExchanger:
    """Device halo exchange: pack kernel -> NCCL send/recv -> unpack kernel, one
    grouped P2P batch per exchange round covering h, u and all tracers."""

    def __init__(self, mesh, lists, ntracers):
        import torch
        self.torch = torch
        self.mesh, self.lib, self.T = mesh, mesh.lib, ntracers
        self.h = lists
        self.nbrs = list(lists["neighbors"])
        nn = len(self.nbrs)
        flat = flatten_lists(lists)
        self._keep = []
        d = capi.HaloDesc(MyTask=int(mesh.L.MyTask), NNeighbors=nn)
        nb = np.array(self.nbrs, np.int32)
        self._keep.append(nb)
        d.NeighborTasks = nb.ctypes.data_as(capi.c_int_p)
        self.counts = {}
        for e, kind in enumerate(KINDS):
            so, si = flat[("send", kind)]
            ro, ri = flat[("recv", kind)]
            self._keep += [so, si, ro, ri]
            d.SendOffsets[e], d.SendIdx[e] = so.ctypes.data_as(capi.c_int_p), si.ctypes.data_as(capi.c_int_p)
            d.RecvOffsets[e], d.RecvIdx[e] = ro.ctypes.data_as(capi.c_int_p), ri.ctypes.data_as(capi.c_int_p)
            self.counts[e] = (so.copy(), ro.copy())
        self.handle = C.c_void_p()
        rc = self.lib.omega_halo_create(mesh.h, C.byref(d), C.byref(self.handle))
        if rc:
            raise RuntimeError("omega_halo_create: " + self.lib.omega_last_error().decode())
        K = mesh.K
        self.bufs = {}
        for key, e, outer in (("h", 0, 1), ("u", 1, 1), ("tr", 0, max(ntracers, 1))):
            so, ro = self.counts[e]
            self.bufs[key] = (torch.empty(int(so[-1]) * outer * K, dtype=torch.float64, device=mesh.device),
                              torch.empty(int(ro[-1]) * outer * K, dtype=torch.float64, device=mesh.device))
        self.bytes_per_round = sum(b[0].numel() * 8 for b in self.bufs.values())

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def pack(self, ptrs):
        K = self.mesh.K
        for key, e, outer, ptr in ptrs:
            sb, _ = self.bufs[key]
            rc = self.lib.omega_halo_pack(self.handle, e, outer, K, ptr, C.c_void_p(sb.data_ptr()), self._stream())
            assert rc == 0, self.lib.omega_last_error()

    def unpack(self, ptrs):
        K = self.mesh.K
        for key, e, outer, ptr in ptrs:
            _, rb = self.bufs[key]
            rc = self.lib.omega_halo_unpack(self.handle, e, outer, K, C.c_void_p(rb.data_ptr()), ptr, self._stream())
            assert rc == 0, self.lib.omega_last_error()

    def segments(self, ptrs):
        """[(neighbour task, send slice, recv slice)] for every array of the round."""
        K = self.mesh.K
        out = []
        for key, e, outer, _ in ptrs:
            sb, rb = self.bufs[key]
            so, ro = self.counts[e]
            for i, nbr in enumerate(self.nbrs):
                s0, s1 = int(so[i]) * outer * K, int(so[i + 1]) * outer * K
                r0, r1 = int(ro[i]) * outer * K, int(ro[i + 1]) * outer * K
                out.append((nbr, sb[s0:s1], rb[r0:r1]))
        return out

    def exchange_round(self, ptrs):
        """ptrs: [(key, elem kind, nouter, device pointer)] -- all arrays of one exchange point."""
        import torch.distributed as dist
        self.pack(ptrs)
        ops = []
        for nbr, s, r in self.segments(ptrs):
            if s.numel():
                ops.append(dist.P2POp(dist.isend, s, nbr))
            if r.numel():
                ops.append(dist.P2POp(dist.irecv, r, nbr))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        self.unpack(ptrs)

    def exchange_list(self, st, state, stepper, phase):
        ptrs = (C.c_void_p * 3)()
        elems, nouter, n = (C.c_int * 3)(), (C.c_int * 3)(), C.c_int()
        rc = self.lib.omega_stepper_exchange_list(st.h, stepper, C.byref(state.c), phase, ptrs, elems, nouter, C.byref(n))
        if rc:
            raise RuntimeError("omega_stepper_exchange_list: " + self.lib.omega_last_error().decode())
        keys = ("h", "u", "tr")
        return [(keys[i], int(elems[i]), int(nouter[i]), C.c_void_p(ptrs[i])) for i in range(n.value)]

    def do_step(self, st, state, stepper, coeffs, fused=True, compat=True):
        """TimeStepper::doStep with the halo exchanges of RungeKutta4Stepper.cpp:108-114
        and OceanState/Tracers::updateTimeLevels."""
        nph = 2 if stepper == capi.RK4 else 1
        for p in range(nph):
            st.stepPhase(state, stepper, coeffs, p, fused=fused, compat=compat)
            self.exchange_round(self.exchange_list(st, state, stepper, p))
        self.lib.omega_state_rotate(C.byref(state.c))

    def close(self):
        if self.handle:
            self.lib.omega_halo_destroy(self.handle)
            self.handle = None
