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

Transport on the GPU: the library itself (csrc/halo.cu: pack kernel, NCCL
send/recv over NVLink, unpack kernel, inside omega_stepper_do_step);
torch.distributed only broadcasts the NCCL unique id.  The numpy versions
below and the gloo exchanger serve the CPU tests of the host logic.
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


def local_problem(mesh, world, rank, nlev, halo_width=decomp.HALO_WIDTH, method=0, use_lib=True):
    """(local mesh of `rank`, its halo lists).  Every rank derives the same partition deterministically and
    numbers only its own partition and its neighbours' (compiled Decomp, csrc/decomp.cpp); no communication.
    use_lib=False: the numpy restatement, which builds every rank's mesh (tests, emulation)."""
    if use_lib and world > 1:
        task, _ = decomp.partition_cells_lib(mesh, world, method)
        return decomp.build_local_lib(mesh, task, rank, nlev, halo_width)
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
    """Device halo exchange through the library (csrc/halo.cu): one pack launch, one grouped NCCL
    send/recv per neighbour and one unpack launch per exchange round (h, u and all tracers), run by
    omega_stepper_do_step itself and overlapped with the interior tiles.  torch.distributed only
    bootstraps the communicator (broadcast of the NCCL unique id)."""

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
        self.bytes_per_round = 8 * K * (int(self.counts[0][0][-1]) * (1 + ntracers) + int(self.counts[1][0][-1]))

    def _chk(self, rc, what):
        if rc:
            raise RuntimeError(what + ": " + self.lib.omega_last_error().decode())

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream().cuda_stream)

    def init_comm(self, world, rank):
        """NCCL communicator over the ranks of torch.distributed's default group (one rank per GPU)."""
        import torch.distributed as dist
        ident = (C.c_char * 128)()
        if rank == 0:
            self._chk(self.lib.omega_halo_nccl_unique_id(ident), "omega_halo_nccl_unique_id")
        dev = self.mesh.device if dist.get_backend() == "nccl" else "cpu"
        t = self.torch.tensor(list(bytes(ident)), dtype=self.torch.uint8, device=dev)
        dist.broadcast(t, 0)
        ident = (C.c_char * 128).from_buffer_copy(bytes(t.cpu().tolist()))
        self._chk(self.lib.omega_halo_comm_init(self.handle, ident, world, rank), "omega_halo_comm_init")

    def attach(self, st):
        self._chk(self.lib.omega_stepper_attach_halo(st.h, self.handle), "omega_stepper_attach_halo")

    def tile_classes(self):
        a, b, c = C.c_int(), C.c_int(), C.c_int()
        self.lib.omega_halo_tile_classes(self.handle, C.byref(a), C.byref(b), C.byref(c))
        return a.value, b.value, c.value

    # -- one round, cut into its device halves (tests: N ranks emulated on one GPU) --------------
    @staticmethod
    def _round(ptrs):
        n = len(ptrs)
        P = (C.c_void_p * 3)(*[p[3] for p in ptrs])
        E = (C.c_int * 3)(*[p[1] for p in ptrs])
        O = (C.c_int * 3)(*[p[2] for p in ptrs])
        return P, E, O, n

    def pack_round(self, ptrs):
        """-> (send buffer address, [send counts per neighbour], [recv counts per neighbour]) in doubles"""
        P, E, O, n = self._round(ptrs)
        buf = C.c_void_p()
        self._chk(self.lib.omega_halo_pack_round(self.handle, P, E, O, n, C.byref(buf), self._stream()), "omega_halo_pack_round")
        nn = len(self.nbrs)
        sc, rcv = (C.c_longlong * max(nn, 1))(), (C.c_longlong * max(nn, 1))()
        self._chk(self.lib.omega_halo_round_counts(self.handle, E, O, n, sc, rcv), "omega_halo_round_counts")
        return buf.value, list(sc)[:nn], list(rcv)[:nn]

    def recv_buffer(self, ptrs):
        P, E, O, n = self._round(ptrs)
        buf = C.c_void_p()
        self._chk(self.lib.omega_halo_recv_buffer(self.handle, E, O, n, C.byref(buf)), "omega_halo_recv_buffer")
        return buf.value

    def unpack_round(self, ptrs):
        P, E, O, n = self._round(ptrs)
        self._chk(self.lib.omega_halo_unpack_round(self.handle, P, E, O, n, self._stream()), "omega_halo_unpack_round")

    def exchange_round(self, ptrs):
        """ptrs: [(key, elem kind, nouter, device pointer)] -- all arrays of one exchange point (NCCL)."""
        P, E, O, n = self._round(ptrs)
        self._chk(self.lib.omega_halo_exchange_round(self.handle, P, E, O, n, self._stream()), "omega_halo_exchange_round")

    def exchange_list(self, st, state, stepper, phase):
        ptrs = (C.c_void_p * 3)()
        elems, nouter, n = (C.c_int * 3)(), (C.c_int * 3)(), C.c_int()
        rc = self.lib.omega_stepper_exchange_list(st.h, stepper, C.byref(state.c), phase, ptrs, elems, nouter, C.byref(n))
        if rc:
            raise RuntimeError("omega_stepper_exchange_list: " + self.lib.omega_last_error().decode())
        keys = ("h", "u", "tr")
        return [(keys[i], int(elems[i]), int(nouter[i]), C.c_void_p(ptrs[i])) for i in range(n.value)]

    def do_step(self, st, state, stepper, coeffs, fused=True, compat=True):
        """TimeStepper::doStep including the halo exchanges of RungeKutta4Stepper.cpp:108-114,130-131 and
        OceanState/Tracers::updateTimeLevels: one library call."""
        self._chk(self.lib.omega_stepper_do_step(st.h, stepper, C.byref(state.c), C.byref(coeffs), self.handle,
                                                 int(fused), int(compat), self._stream()), "omega_stepper_do_step")

    def set_profiling(self, on):
        self.lib.omega_halo_set_profiling(self.handle, int(on))

    def profile(self):
        ms, rounds, nbytes = C.c_double(), C.c_longlong(), C.c_longlong()
        self._chk(self.lib.omega_halo_get_profile(self.handle, C.byref(ms), C.byref(rounds), C.byref(nbytes)), "omega_halo_get_profile")
        return ms.value, rounds.value, nbytes.value

    def close(self):
        if self.handle:
            self.lib.omega_halo_destroy(self.handle)
            self.handle = None
