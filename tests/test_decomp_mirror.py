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


# ---- the whole host side in C++ from the global mesh: tests/cpp/global_mesh_test.cpp ----
EXE_G = os.path.join(ROOT, "tests", "cpp", "global_mesh_test")
GEOMETRY = (("AreaCell", "areaCell"), ("BottomDepth", "bottomDepth"), ("XCell", "xCell"), ("YCell", "yCell"), ("ZCell", "zCell"),
            ("DcEdge", "dcEdge"), ("DvEdge", "dvEdge"), ("AngleEdge", "angleEdge"), ("FEdge", "fEdge"), ("XEdge", "xEdge"),
            ("YEdge", "yEdge"), ("WeightsOnEdge", "weightsOnEdge"), ("AreaTriangle", "areaTriangle"), ("FVertex", "fVertex"),
            ("KiteAreasOnVertex", "kiteAreasOnVertex"))


def build_exe_global():
    src = os.path.join(ROOT, "tests", "cpp", "global_mesh_test.cpp")
    lib_dir = os.path.join(ROOT, "omega_b200", "csrc")
    cmd = ["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "omega_b200", "host"),
           "-I/usr/local/cuda/include", src, "-o", EXE_G, "-L" + lib_dir, "-lomega_b200", "-L/usr/local/cuda/lib64", "-lcudart",
           "-Wl,-rpath," + lib_dir, "-Wl,-rpath,/usr/local/cuda/lib64"]
    subprocess.run(cmd, check=True)


def write_global_case(d, level, K, T, nsteps, halo_width):
    """global mesh + geometry + TC5 state in GLOBAL order + config; returns what the checks need"""
    m = cases.sphere_mesh(level)
    L1, cfg, (h, u, tr), dt = cases.sphere_case(level, K, T)
    write_global_mesh(d, m)
    for _, src in GEOMETRY:
        if src not in ("xCell", "yCell", "zCell"):
            np.ascontiguousarray(m[src], np.float64).tofile(os.path.join(d, src + ".bin"))
    cell_of, edge_of = L1.CellID[:-1] - 1, L1.EdgeID[:-1] - 1
    H = np.empty_like(h[:-1]); H[cell_of] = h[:-1]
    U = np.empty_like(u[:-1]); U[edge_of] = u[:-1]
    TR = np.empty_like(tr[:, :-1]); TR[:, cell_of] = tr[:, :-1]
    H.tofile(os.path.join(d, "h.bin")); U.tofile(os.path.join(d, "u.bin")); TR.tofile(os.path.join(d, "tr.bin"))
    with open(os.path.join(d, "cfg.bin"), "wb") as f:
        f.write(bytes(cfg))
    with open(os.path.join(d, "meta.txt"), "w") as f:
        f.write(f"NVertLayers {K}\nNTracers {T}\nNSteps {nsteps}\nHaloWidth {halo_width}\nDt {dt}\n")
    return m, L1, cfg, (h, u, tr), dt


def test_global_mesh_program_local_geometry_and_state(exe, tmp_path):
    """"dry" run, no device: the local geometry and the gathered initial state of every task equal the restatement's"""
    from omega_b200 import analytic
    build_exe_global()
    K, T, world, width = 3, 2, 3, 4
    d = str(tmp_path)
    m, *_ = write_global_case(d, 3, K, T, 1, width)
    locs = decomp.build_all(m, world, K, width)     # recursive coordinate bisection == PartMethodRCB
    for r, L in enumerate(locs):
        res = subprocess.run([EXE_G, d, str(r), str(world), "dry"], capture_output=True, text=True)
        assert res.returncode == 0, res.stderr
        get = lambda n, dt=np.float64: np.fromfile(os.path.join(d, f"r{r}_{n}.bin"), dt)
        assert np.array_equal(get("CellID", np.int32), L.CellID[:-1]) and np.array_equal(get("EdgeID", np.int32), L.EdgeID[:-1])
        for name, _ in GEOMETRY:
            assert np.array_equal(get(name), np.asarray(L[name], np.float64).ravel()), name
        h, u, tr = analytic.tc5(L, T)
        for n, a in (("h0", h), ("u0", u), ("tr0", tr)):
            assert np.array_equal(get(n), a.ravel()), n


def _check_global_run(d, world, m, L1, cfg, state, dt, K, T, nsteps, width):
    from tests.oracle_lib import Oracle
    h, u, tr = state
    o = Oracle(L1, cfg, T)
    o.make_state(h, u, tr)
    for s in range(nsteps):
        o.do_step(capi.RK4, capi.rk_coeffs(capi.RK4, dt, t0=s * dt))
    gh, gu, gt = o.current()
    cell_of, edge_of = L1.CellID[:-1] - 1, L1.EdgeID[:-1] - 1
    H = np.empty_like(gh[:-1]); H[cell_of] = gh[:-1]
    U = np.empty_like(gu[:-1]); U[edge_of] = gu[:-1]
    TR = np.empty_like(gt[:, :-1]); TR[:, cell_of] = gt[:, :-1]
    o.close()
    seen = np.zeros(m.nCells, bool)
    for r in range(world):
        get = lambda n, dt_=np.float64: np.fromfile(os.path.join(d, f"r{r}_{n}.bin"), dt_)
        ci, ei = get("CellID", np.int32) - 1, get("EdgeID", np.int32) - 1
        lh, lu, lt = get("h").reshape(-1, K), get("u").reshape(-1, K), get("tr").reshape(T, -1, K)
        for a, b in ((lh[:ci.size], H[ci]), (lu[:ei.size], U[ei]), (lt[:, :ci.size], TR[:, ci])):
            assert cases.rel_l2(a, b) <= 1e-10, (r, cases.rel_l2(a, b))
        seen[ci] = True
    assert seen.all()


@pytest.mark.gpu
def test_global_mesh_program_one_task(tmp_path):
    """global mesh -> Decomp mirror -> HorzMesh -> steppers, all in C++, one task on one GPU, against the oracle"""
    build_exe_global()
    K, T, nsteps, width = 16, 2, 3, 3
    d = str(tmp_path)
    m, L1, cfg, state, dt = write_global_case(d, 3, K, T, nsteps, width)
    subprocess.run([EXE_G, d, "0", "1"], check=True, timeout=300)
    _check_global_run(d, 1, m, L1, cfg, state, dt, K, T, nsteps, width)


@pytest.mark.gpu
def test_global_mesh_program_two_tasks(tmp_path):
    """the same with two C++ host processes, one GPU each: each numbers its own partition from the global mesh, the
    Halo is built from the Decomp mirror's lists, the exchange rounds run inside doStep over NCCL"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    build_exe_global()
    K, T, nsteps, width, world = 16, 2, 3, 4, 2
    d = str(tmp_path)
    m, L1, cfg, state, dt = write_global_case(d, 4, K, T, nsteps, width)
    procs = [subprocess.Popen([EXE_G, d, str(r), str(world)]) for r in range(world)]
    for p in procs:
        assert p.wait(timeout=300) == 0
    _check_global_run(d, world, m, L1, cfg, state, dt, K, T, nsteps, width)
