"""Pins the multi-task local numbering against the REFERENCE ITSELF: src/base/Decomp.cpp, compiled unmodified over
oracle/ref_decomp/decomp_shim.h and run as N tasks = N threads of one process (Broadcast = thread rendezvous, the mesh
file = memory, METIS = the partition handed in).  Every count, halo-layer count, global ID, owner address and local
connectivity entry of every task must equal both the numpy restatement (omega_b200/decomp.py) and the compiled
Decomp of the library (csrc/decomp.cpp) -- including the boundary-slot entries.  This is what SURVEY.md 3.5 P4 (reverse
first-halo edge order etc.) and VERDICT r1 called parity unpinned."""
import ctypes as C
import os

import numpy as np
import pytest

from omega_b200 import decomp
from tests import cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "libomega_ref_decomp.so")
ip = C.POINTER(C.c_int)
P = lambda a: a.ctypes.data_as(ip)
NAMES = {0: "CellID", 1: "EdgeID", 2: "VertexID", 3: "CellLoc", 4: "EdgeLoc", 5: "VertexLoc", 10: "CellsOnCell",
         11: "EdgesOnCell", 12: "VerticesOnCell", 13: "CellsOnEdge", 14: "VerticesOnEdge", 15: "EdgesOnEdge",
         16: "CellsOnVertex", 17: "EdgesOnVertex", 18: "NEdgesOnCell", 19: "NEdgesOnEdge"}


def _mesh(kind):
    if kind.startswith("sphere"):
        return cases.sphere_mesh(int(kind[-1]))
    if kind.startswith("hept"):
        return cases.hept_mesh(int(kind[-1]))      # maxEdges = 7, maxEdges2 = 14
    if kind.startswith("land"):
        return cases.land_mesh(int(kind[-1]))      # coastline: -1 entries in every connectivity array
    return cases.plane_mesh(16, 16, 1000.0)


def ref_lib():
    if not os.path.exists(REF):
        mk = os.path.join(ROOT, "oracle", "ref_decomp")
        if not os.path.isdir("/root/reference/components/omega/src"):
            pytest.skip("oracle/_ref/libomega_ref_decomp.so not built and /root/reference absent")
        import subprocess
        subprocess.run(["make", "-s", "-C", mk], check=True)
    return C.CDLL(REF)


def run_reference(lib, m, ntasks, hw, task):
    """the reference's Decomp on ntasks thread-tasks; file conventions: 1-based IDs, 0 = none"""
    a = {n: np.ascontiguousarray(np.asarray(m[n]) + 1, np.int32) for n in (
        "cellsOnCell", "edgesOnCell", "verticesOnCell", "cellsOnEdge", "edgesOnEdge", "verticesOnEdge", "cellsOnVertex", "edgesOnVertex")}
    assert m.maxEdges2 == 2 * m.maxEdges
    task = np.ascontiguousarray(task, np.int32)
    rc = lib.ref_decomp_run(ntasks, hw, m.nCells, m.nEdges, m.nVertices, m.maxEdges, m.vertexDegree, P(a["cellsOnCell"]),
                            P(a["edgesOnCell"]), P(a["verticesOnCell"]), P(a["cellsOnEdge"]), P(a["edgesOnEdge"]),
                            P(a["verticesOnEdge"]), P(a["cellsOnVertex"]), P(a["edgesOnVertex"]), P(task))
    assert rc == 0
    out = []
    for t in range(ntasks):
        cnt = np.zeros(3 * (2 + hw), np.int32)
        lib.ref_decomp_counts(t, P(cnt))
        d = {"counts": cnt.reshape(3, 2 + hw)}
        for w, nm in NAMES.items():
            n = lib.ref_decomp_get(t, w, None)
            v = np.zeros(n, np.int32)
            lib.ref_decomp_get(t, w, P(v))
            d[nm] = v
        out.append(d)
    return out


@pytest.mark.parametrize("kind,ntasks,hw", [("sphere3", 2, 3), ("sphere3", 3, 3), ("sphere4", 5, 4), ("sphere4", 8, 3),
                                            ("sphere3", 4, 1), ("plane", 4, 3), ("plane", 6, 2), ("sphere3", 1, 3),
                                            ("land3", 1, 3), ("land3", 3, 3), ("land4", 4, 4), ("hept3", 3, 3),
                                            # small partitions: halos that reach beyond the neighbours' neighbours
                                            ("sphere2", 8, 3), ("sphere3", 16, 3), ("sphere3", 24, 5), ("land3", 7, 2),
                                            # periodic plane: halos that wrap around and meet themselves
                                            ("plane", 2, 5), ("plane", 3, 4)])
def test_local_numbering_equals_the_reference_decomp(kind, ntasks, hw):
    lib = ref_lib()
    m = _mesh(kind)
    task = decomp.partition_cells(m, ntasks) if ntasks > 1 else np.zeros(m.nCells, np.int32)
    ref = run_reference(lib, m, ntasks, hw, task)
    locs = decomp.build_all(m, ntasks, 3, hw, cell_task=task)
    have_lib = os.path.exists(os.path.join(ROOT, "omega_b200", "csrc", "libomega_b200.so"))
    for t in range(ntasks):
        R = ref[t]
        ours = [("numpy restatement", locs[t])]
        if have_lib:
            ours.append(("csrc/decomp.cpp", decomp.build_local_lib(m, np.ascontiguousarray(task, np.int32), t, 3, hw)[0]))
        for label, L in ours:
            for k, nm in enumerate(("Cells", "Edges", "Vertices")):
                assert int(L["N%sOwned" % nm]) == R["counts"][k, 0] and int(L["N%sAll" % nm]) == R["counts"][k, 1], (label, t, nm)
                assert np.array_equal(np.asarray(L["N%sHalo" % nm]), R["counts"][k, 2:]), (label, t, nm, "halo layers")
            for nm in NAMES.values():
                a, b = R[nm], np.asarray(L[nm]).astype(np.int32).ravel()
                assert a.size == b.size and np.array_equal(a, b), (label, t, nm, int(np.argmax(a != b)) if a.size == b.size else "size")


@pytest.mark.parametrize("kind,ntasks,hw", [("sphere3", 2, 3), ("sphere3", 3, 3), ("sphere4", 5, 4), ("sphere4", 8, 3),
                                            ("sphere3", 4, 1), ("plane", 4, 3), ("plane", 6, 2), ("land3", 3, 3), ("land4", 4, 4), ("hept3", 3, 3),
                                            ("sphere2", 8, 3), ("sphere3", 16, 3), ("sphere3", 24, 5), ("land3", 7, 2), ("plane", 2, 5), ("plane", 3, 4)])
def test_halo_exchange_lists_equal_the_reference_halo(kind, ntasks, hw):
    """src/base/Halo.cpp (constructor: neighbours, send / receive lists per index space and halo layer), compiled
    unmodified and run on the reference Decomp of every thread-task over a thread MPI (oracle/ref_decomp/mpi.h:
    MPI_Alltoall, Isend / Irecv / Waitall).  Neighbour tasks, every list's indices in order, and the per-layer counts
    equal the lists the library is given (parallel.build_halo_lists, and omega_decomp_halo_lists of csrc/decomp.cpp)."""
    from omega_b200 import parallel
    lib = ref_lib()
    m = _mesh(kind)
    task = decomp.partition_cells(m, ntasks)
    run_reference(lib, m, ntasks, hw, task)
    locs = decomp.build_all(m, ntasks, 3, hw, cell_task=task)
    lists = parallel.build_halo_lists(locs)
    have_lib = os.path.exists(os.path.join(ROOT, "omega_b200", "csrc", "libomega_b200.so"))
    for t in range(ntasks):
        nn = lib.ref_halo_neighbors(t, None)
        nb = np.zeros(max(nn, 1), np.int32)
        lib.ref_halo_neighbors(t, P(nb))
        assert nb[:nn].tolist() == lists[t]["neighbors"], (t, "neighbours")
        ours = [("numpy", lists[t])]
        if have_lib:
            ours.append(("csrc/decomp.cpp", decomp.build_local_lib(m, np.ascontiguousarray(task, np.int32), t, 3, hw)[1]))
        L = locs[t]
        for i, nbr in enumerate(nb[:nn].tolist()):
            for e, k in enumerate(parallel.KINDS):
                for d, dname in ((0, "send"), (1, "recv")):
                    n = lib.ref_halo_list(t, i, e, d, None, None)
                    nl = hw if k == "Cell" else hw + 1       # edges and vertices: one layer more (Halo.cpp:469-499)
                    idx, nh = np.zeros(max(n, 1), np.int32), np.zeros(nl, np.int32)
                    lib.ref_halo_list(t, i, e, d, P(idx), P(nh))
                    assert int(nh.sum()) == n
                    for label, o in ours:
                        mine = o[dname][k].get(nbr, np.zeros(0, np.int32))
                        assert np.array_equal(idx[:n], np.asarray(mine)), (label, t, nbr, k, dname)
                    if d == 1 and n:
                        # receive lists are grouped by halo layer, ascending inside (Halo.cpp:455-600): the per-layer
                        # counts are the elements of that owner in each layer of the local numbering
                        kk = "Vertices" if k == "Vertex" else k + "s"
                        bounds = np.concatenate(([int(L["N%sOwned" % kk])], np.asarray(L["N%sHalo" % kk]), [int(L["N%sAll" % kk])]))
                        want = [int(((idx[:n] >= bounds[h]) & (idx[:n] < bounds[h + 1])).sum()) for h in range(nl)]
                        assert nh.tolist() == want, (t, nbr, k, "layer counts")


def _ref_exchange(lib, locs, kind_idx, nouter, ninner, arrays):
    ptrs = (C.POINTER(C.c_double) * len(arrays))(*[a.ctypes.data_as(C.POINTER(C.c_double)) for a in arrays])
    assert lib.ref_halo_exchange(kind_idx, nouter, ninner, ptrs) == 0


@pytest.mark.parametrize("kind,ntasks,hw", [("sphere3", 3, 3), ("sphere4", 5, 4), ("plane", 4, 2), ("land3", 3, 3), ("sphere2", 8, 3)])
def test_reference_exchange_and_message_layout(kind, ntasks, hw):
    """Halo::exchangeFullArrayHalo of the reference itself (Halo.h:768-915) on the thread-tasks: (1) its own HaloTest
    (test/base/HaloTest.cpp:41-99) passes on them -- arrays holding f(global ID) in the owned entries equal the fully
    initialised arrays after the exchange, 1-D, 2-D and 3-D, cells, edges, vertices; (2) the message it packs for a
    neighbour is [outer][position in the send list][inner], the layout csrc/halo.cu restates
    (tests/test_halo_any_type.py::test_pack_layout_equals_the_reference_message checks the kernel against it)."""
    from omega_b200 import parallel
    lib = ref_lib()
    m = _mesh(kind)
    task = decomp.partition_cells(m, ntasks)
    run_reference(lib, m, ntasks, hw, task)
    locs = decomp.build_all(m, ntasks, 3, hw, cell_task=task)
    lists = parallel.build_halo_lists(locs)
    for e, (k, idn, kk) in enumerate((("Cell", "CellID", "Cells"), ("Edge", "EdgeID", "Edges"), ("Vertex", "VertexID", "Vertices"))):
        for nouter, ninner in ((0, 0), (0, 5), (2, 3)):
            full, part = [], []
            for L in locs:
                n, nall, nown = int(L["N%sSize" % kk]), int(L["N%sAll" % kk]), int(L["N%sOwned" % kk])
                ids = np.asarray(L[idn], np.float64)[:n]
                o = np.arange(max(nouter, 1)).reshape(-1, 1, 1)
                i = np.arange(max(ninner, 1)).reshape(1, 1, -1)
                f = (ids.reshape(1, n, 1) * 7 + o) * 5 + i
                f[:, nall:] = 0.0
                p = f.copy()
                p[:, nown:nall] = -1.0
                shape = ((nouter,) if nouter else ()) + (n,) + ((ninner,) if ninner else ())
                full.append(np.ascontiguousarray(f.reshape(shape)))
                part.append(np.ascontiguousarray(p.reshape(shape)))
            _ref_exchange(lib, locs, e, nouter, ninner, part)
            for t in range(ntasks):
                assert np.array_equal(part[t], full[t]), (k, nouter, ninner, t)
                for j, nbr in enumerate(lists[t]["neighbors"]):
                    idx = lists[t]["send"][k].get(nbr, np.zeros(0, np.int32))
                    cnt = idx.size * max(nouter, 1) * max(ninner, 1)
                    buf = np.zeros(max(cnt, 1))
                    have = lib.ref_halo_send_buffer(t, j, cnt, buf.ctypes.data_as(C.POINTER(C.c_double)))
                    assert have >= cnt
                    a = full[t].reshape(max(nouter, 1), -1, max(ninner, 1))
                    assert np.array_equal(buf[:cnt], a[:, idx, :].ravel()), (k, nouter, ninner, t, nbr, "message layout")


@pytest.mark.parametrize("kind,ntasks,hw", [("sphere3", 1, 3), ("sphere3", 3, 3), ("sphere4", 5, 4), ("plane", 4, 2),
                                            ("land3", 1, 3), ("land4", 4, 3), ("hept3", 2, 3)])
def test_horzmesh_local_geometry_and_derived_fields_equal_the_reference(kind, ntasks, hw):
    """src/ocn/HorzMesh.cpp compiled unmodified and constructed on every task's reference Decomp (reads its local
    geometry through the IO decompositions it builds from the global IDs; computeEdgeSign, setMasks, setMeshScaling:
    HorzMesh.cpp:527-647).  The local geometry arrays equal those of omega_b200/decomp.py entry for entry, and
    EdgeSignOnCell / EdgeSignOnVertex / EdgeMask equal what the C oracle derives (the same arrays the library's
    omega_mesh_derived is held to bit for bit on the GPU)."""
    from omega_b200 import capi
    from tests.oracle_lib import Oracle
    lib = ref_lib()
    m = _mesh(kind)
    task = decomp.partition_cells(m, ntasks) if ntasks > 1 else np.zeros(m.nCells, np.int32)
    run_reference(lib, m, ntasks, hw, task)
    keep = []
    for nm in ("xCell", "yCell", "zCell", "lonCell", "latCell", "xEdge", "yEdge", "zEdge", "lonEdge", "latEdge", "xVertex", "yVertex",
               "zVertex", "lonVertex", "latVertex", "bottomDepth", "areaCell", "areaTriangle", "dvEdge", "dcEdge", "angleEdge",
               "meshDensity", "fCell", "fVertex", "fEdge", "kiteAreasOnVertex", "weightsOnEdge"):
        a = np.ascontiguousarray(m[nm], np.float64) if nm in m else None
        keep.append(a)
        lib.ref_mesh_set_r8(nm.encode(), a.ctypes.data_as(C.POINTER(C.c_double)) if a is not None else None)
    K = 3
    assert lib.ref_horzmesh_run(K) == 0
    locs = decomp.build_all(m, ntasks, K, hw, cell_task=task)
    dp = C.POINTER(C.c_double)

    def get(t, w):
        n = lib.ref_horzmesh_get(t, w, None)
        v = np.zeros(n)
        lib.ref_horzmesh_get(t, w, v.ctypes.data_as(dp))
        return v

    geo = {3: "AreaCell", 4: "DcEdge", 5: "DvEdge", 6: "AngleEdge", 7: "FEdge", 8: "FVertex", 9: "AreaTriangle", 10: "BottomDepth",
           11: "WeightsOnEdge", 12: "KiteAreasOnVertex", 15: "XCell", 16: "YCell", 17: "XEdge", 18: "YEdge"}
    cfg = capi.TendConfig.default()
    for t in range(ntasks):
        L = locs[t]
        for w, nm in geo.items():
            a, b = get(t, w), np.asarray(L[nm], np.float64)
            rows = b.shape[0] - 1                                  # every local element; the boundary slot apart
            width = 1 if b.ndim == 1 else b.shape[1]
            assert a.size == b.size, (t, nm)
            assert np.array_equal(a.reshape(-1, width)[:rows], b.reshape(-1, width)[:rows]), (t, nm)
        o = Oracle(L, cfg, 0)
        esc, esv, mask = o.derived()
        assert np.array_equal(get(t, 0).reshape(esc.shape), esc), (t, "EdgeSignOnCell")
        assert np.array_equal(get(t, 1).reshape(esv.shape), esv), (t, "EdgeSignOnVertex")
        em = get(t, 2).reshape(mask.shape[0], K)
        for k in range(K):
            assert np.array_equal(em[:, k], mask), (t, "EdgeMask", k)
        assert np.all(get(t, 13)[:int(L.NEdgesAll)] == 1.0) and np.all(get(t, 14)[:int(L.NEdgesAll)] == 1.0)   # HorzMesh.cpp:604-626
        o.close()
