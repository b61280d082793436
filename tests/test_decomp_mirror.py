"""The C++ Decomp mirror (omega_b200/host/OmegaB200Decomp.h: the reference's public members, create / get / getDefault /
erase / clear, valid*ID) driven by tests/cpp/decomp_mirror_test.cpp as test/base/DecompTest.cpp drives the reference's
class.  Every public array is compared with the numpy restatement (omega_b200/decomp.py), which
tests/test_decomp_vs_reference.py pins entry by entry to the reference's own Decomp.cpp and Halo.cpp.  Host code only."""
import os
import subprocess

import numpy as np
import pytest

from omega_b200 import capi, decomp, parallel
from tests import cases

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "decomp_mirror_test")
KINDS = ("Cell", "Edge", "Vertex")


def build_exe():
    src = os.path.join(ROOT, "tests", "cpp", "decomp_mirror_test.cpp")
    lib_dir = os.path.join(ROOT, "omega_b200", "csrc")
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "omega_b200", "host"),
           src, "-o", EXE, "-L" + lib_dir, "-lomega_b200", "-L/usr/local/cuda/lib64", "-lcudart",
           "-Wl,-rpath," + lib_dir, "-Wl,-rpath,/usr/local/cuda/lib64"]
    subprocess.run(cmd, check=True)


@pytest.fixture(scope="module")
def exe():
    if not os.path.exists(capi.LIB_PATH):
        pytest.skip("libomega_b200.so not built")
    build_exe()
    return EXE


def write_global_mesh(d, m):
    os.makedirs(d, exist_ok=True)
    np.array([m.nCells, m.nEdges, m.nVertices, m.maxEdges, m.maxEdges2, m.vertexDegree], np.int32).tofile(os.path.join(d, "dims.bin"))
    for n in ("cellsOnCell", "edgesOnCell", "verticesOnCell", "cellsOnEdge", "verticesOnEdge", "edgesOnEdge", "cellsOnVertex",
              "edgesOnVertex"):
        np.ascontiguousarray(m[n], np.int32).tofile(os.path.join(d, n + ".bin"))
    for n in ("xCell", "yCell", "zCell"):
        np.ascontiguousarray(m[n], np.float64).tofile(os.path.join(d, n + ".bin"))


def out(d, name):
    return np.fromfile(os.path.join(d, "out_" + name + ".bin"), np.int32)


@pytest.mark.parametrize("mesh_kind,ntasks,halo_width,method", [
    ("sphere", 1, 3, 0), ("sphere", 3, 3, 0), ("sphere", 4, 4, 1), ("plane", 2, 3, 0)])
def test_decomp_mirror_members_equal_the_pinned_numbering(exe, tmp_path, mesh_kind, ntasks, halo_width, method):
    m = cases.sphere_mesh(3) if mesh_kind == "sphere" else cases.plane_mesh(12, 12, 1.0e4)
    task_ref, cut_ref = (decomp.partition_cells_lib(m, ntasks, method) if ntasks > 1 else (np.zeros(m.nCells, np.int32), 0))
    locs = decomp.build_all(m, ntasks, 1, halo_width, cell_task=task_ref)
    lists = parallel.build_halo_lists(locs)
    owned_sum = 0
    for t in range(ntasks):
        d = str(tmp_path / f"t{t}")
        write_global_mesh(d, m)
        r = subprocess.run([exe, d, str(ntasks), str(t), str(halo_width), str(method)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        meta = dict(l.split() for l in open(os.path.join(d, "out_meta.txt")))
        L = locs[t]
        assert np.array_equal(out(d, "CellTask"), task_ref)
        assert int(meta["EdgeCut"]) == int(cut_ref)
        for nm in ("Cells", "Edges", "Vertices"):
            assert int(meta[f"N{nm}Owned"]) == L[f"N{nm}Owned"] and int(meta[f"N{nm}All"]) == L[f"N{nm}All"]
            assert np.array_equal(out(d, f"N{nm}HaloH"), L[f"N{nm}Halo"])
        for kind in KINDS:
            assert np.array_equal(out(d, kind + "IDH"), L[kind + "ID"]), kind
            assert np.array_equal(out(d, kind + "LocH"), L[kind + "Loc"].ravel()), kind
        for n in ("CellsOnCell", "EdgesOnCell", "VerticesOnCell", "CellsOnEdge", "VerticesOnEdge", "EdgesOnEdge", "CellsOnVertex",
                  "EdgesOnVertex", "NEdgesOnCell", "NEdgesOnEdge"):
            assert np.array_equal(out(d, n + "H"), np.asarray(L[n]).ravel()), n
        h = lists[t]
        nbrs = out(d, "NeighborTasks")
        assert list(nbrs) == h["neighbors"]
        for kind in KINDS:
            for way, off, idx in (("send", "SendOffsets", "SendIdx"), ("recv", "RecvOffsets", "RecvIdx")):
                o, i = out(d, off + kind), out(d, idx + kind)
                for j, nb in enumerate(nbrs):
                    want = h[way][kind].get(int(nb), np.zeros(0, np.int32))
                    assert np.array_equal(i[o[j]:o[j + 1]], want), (kind, way, nb)
        owned_sum += int(meta["OwnedIDSum"])
    # test/base/DecompTest.cpp:126-136: the owned IDs of all tasks are 1..NCells once each
    assert owned_sum == m.nCells * (m.nCells + 1) // 2
