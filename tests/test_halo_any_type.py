"""Halo exchange of arrays of any element type and rank, after test/base/HaloTest.cpp:41-99 of the reference: an
array holds global IDs in its owned entries only, is exchanged, and must then equal the array initialised from the
global IDs everywhere (halo rows included).  I4 / I8 / R4 / R8, one to five dimensions, cells, edges, vertices.
N ranks are emulated on one GPU: the library packs and unpacks (omega_halo_pack_any / unpack_any), device-to-device
copies stand in for NCCL (tests/nccl_check.py runs omega_halo_exchange_any itself on a multi-GPU box)."""
import ctypes as C

import numpy as np
import pytest

from omega_b200 import capi
from tests import cases

pytestmark = pytest.mark.gpu

SHAPES = [((), ()), ((), (5,)), ((2,), (5,)), ((2, 3), (4,)), ((2, 2, 2), (3,))]     # (dims in front, dims behind) the element dim
DTYPES = [np.int32, np.int64, np.float32, np.float64]


@pytest.mark.parametrize("ntasks", [2, 3])
def test_exchange_any_type_any_rank(ntasks):
    import torch
    from omega_b200 import decomp, model, parallel
    mesh_g = cases.sphere_mesh(3)
    locs = decomp.build_all(mesh_g, ntasks, 4, 2)
    lists = parallel.build_halo_lists(locs)
    ranks = []
    for L, hl in zip(locs, lists):
        m = model.HorzMesh(L, reorder=True)
        ranks.append((L, m, parallel.HaloExchanger(m, hl, 0)))
    lib = ranks[0][1].lib
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    tdt = {np.int32: torch.int32, np.int64: torch.int64, np.float32: torch.float32, np.float64: torch.float64}
    for kind, idname, nall, nown in ((capi.ON_CELL, "CellID", "NCellsAll", "NCellsOwned"), (capi.ON_EDGE, "EdgeID", "NEdgesAll", "NEdgesOwned"),
                                     (capi.ON_VERTEX, "VertexID", "NVerticesAll", "NVerticesOwned")):
        for dt in DTYPES:
            eb = np.dtype(dt).itemsize
            for front, back in SHAPES:
                nouter, ninner = int(np.prod(front)) if front else 1, int(np.prod(back)) if back else 1
                init, test, bufs = [], [], []
                for L, m, ex in ranks:
                    n = m.NSize[kind]
                    ids = np.asarray(getattr(L, idname), np.int64)[:n]
                    # value = f(global ID, outer index, inner index): distinct per entry, exact in every type
                    o = np.arange(nouter).reshape(nouter, 1, 1)
                    i = np.arange(ninner).reshape(1, 1, ninner)
                    full = ((ids.reshape(1, n, 1) * 7 + o) * 5 + i).astype(dt)
                    full[:, int(getattr(L, nall)):] = 0            # boundary slot
                    part = full.copy()
                    part[:, int(getattr(L, nown)):int(getattr(L, nall))] = dt(-1)      # halo entries unset
                    p, _ = m.perm(kind)
                    init.append(full[:, p])
                    d = torch.from_numpy(np.ascontiguousarray(part[:, p])).to("cuda:0")
                    assert d.dtype == tdt[dt]
                    test.append(d)
                    sc = (C.c_longlong * max(len(ex.nbrs), 1))()
                    rc = (C.c_longlong * max(len(ex.nbrs), 1))()
                    assert lib.omega_halo_buffer_counts(ex.handle, kind, nouter, ninner, sc, rc) == 0
                    sc, rc = list(sc)[:len(ex.nbrs)], list(rc)[:len(ex.nbrs)]
                    sb = torch.zeros(max(sum(sc), 1) * eb, dtype=torch.uint8, device="cuda:0")
                    rb = torch.zeros(max(sum(rc), 1) * eb, dtype=torch.uint8, device="cuda:0")
                    bufs.append((sc, rc, sb, rb))
                    assert lib.omega_halo_pack_any(ex.handle, kind, nouter, ninner, eb, C.c_void_p(d.data_ptr()), C.c_void_p(sb.data_ptr()), stream) == 0, \
                        lib.omega_last_error().decode()
                for t, ((L, m, ex), (sc, rc, sb, rb)) in enumerate(zip(ranks, bufs)):
                    soff = np.concatenate(([0], np.cumsum(sc)))
                    for i2, nbr in enumerate(ex.nbrs):
                        sc2, rc2, sb2, rb2 = bufs[nbr]
                        j = ranks[nbr][2].nbrs.index(t)
                        roff2 = np.concatenate(([0], np.cumsum(rc2)))
                        assert rc2[j] == sc[i2]
                        assert lib.omega_dev_copy(C.c_void_p(rb2.data_ptr() + eb * int(roff2[j])), C.c_void_p(sb.data_ptr() + eb * int(soff[i2])),
                                                  C.c_ulonglong(eb * sc[i2]), stream) == 0
                for (L, m, ex), (sc, rc, sb, rb), d, want in zip(ranks, bufs, test, init):
                    assert lib.omega_halo_unpack_any(ex.handle, kind, nouter, ninner, eb, C.c_void_p(rb.data_ptr()), C.c_void_p(d.data_ptr()), stream) == 0
                    torch.cuda.synchronize()
                    assert np.array_equal(d.cpu().numpy(), want), (kind, dt, front, back)
    assert lib.omega_halo_pack_any(ranks[0][2].handle, 0, 1, 1, 2, C.c_void_p(1), C.c_void_p(1), stream) == 2      # 2-byte elements: rejected
    for L, m, ex in ranks:
        ex.close(); m.close()


def test_pack_layout_equals_the_reference_message():
    """The message the pack kernel builds for a neighbour is the one the reference's Halo packs -- [outer][position in
    the send list][inner] (Halo.h:304-502), pinned on the reference's own code by
    tests/test_decomp_vs_reference.py::test_reference_exchange_and_message_layout."""
    import torch
    from omega_b200 import decomp, model, parallel
    mesh_g = cases.sphere_mesh(3)
    ntasks, K = 3, 5
    locs = decomp.build_all(mesh_g, ntasks, K, 3)
    lists = parallel.build_halo_lists(locs)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for t, (L, hl) in enumerate(zip(locs, lists)):
        m = model.HorzMesh(L, reorder=True)
        ex = parallel.HaloExchanger(m, hl, 0)
        lib = m.lib
        for kind, k in ((capi.ON_CELL, "Cell"), (capi.ON_EDGE, "Edge"), (capi.ON_VERTEX, "Vertex")):
            n = m.NSize[kind]
            for nouter in (1, 2):
                host = np.random.default_rng(5 + t).standard_normal((nouter, n, K))
                p, _ = m.perm(kind)
                d = torch.from_numpy(np.ascontiguousarray(host[:, p])).to("cuda:0")
                sc = (C.c_longlong * max(len(ex.nbrs), 1))()
                rc = (C.c_longlong * max(len(ex.nbrs), 1))()
                assert lib.omega_halo_buffer_counts(ex.handle, kind, nouter, K, sc, rc) == 0
                sc = list(sc)[:len(ex.nbrs)]
                sb = torch.zeros(max(sum(sc), 1), dtype=torch.float64, device="cuda:0")
                assert lib.omega_halo_pack_any(ex.handle, kind, nouter, K, 8, C.c_void_p(d.data_ptr()), C.c_void_p(sb.data_ptr()), stream) == 0
                torch.cuda.synchronize()
                got, off = sb.cpu().numpy(), 0
                for j, nbr in enumerate(ex.nbrs):
                    idx = hl["send"][k].get(nbr, np.zeros(0, np.int32))
                    assert sc[j] == idx.size * nouter * K
                    assert np.array_equal(got[off:off + sc[j]], host[:, idx, :].ravel()), (t, k, nouter, nbr)
                    off += sc[j]
        ex.close(); m.close()
