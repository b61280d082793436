"""File I/O both ways: classic netCDF CDF-2 / CDF-5 writer + reader (omega_b200/ncio.py),
restart files with State + tracers + SimulationTime (src/ocn/OceanInit.cpp:152-177)."""
import os

import numpy as np
import pytest

from omega_b200 import analytic, capi, decomp, meshgen, ncio, parallel
from tests import cases
from tests.oracle_lib import Oracle


def test_cdf2_written_here_is_read_by_scipy_and_vice_versa(tmp_path):
    from scipy.io import netcdf_file
    p = str(tmp_path / "a.nc")
    w = ncio.CDFWriter(p, version=2)
    w.dim("nCells", 7); w.dim("nVertLevels", 3); w.dim("TWO", 2)
    w.att("on_a_sphere", "YES"); w.att("sphere_radius", 6371220.0); w.att("n", np.array([3], np.int32))
    w.variable("layerThickness", "f8", ("nCells", "nVertLevels"), {"units": "m"})
    w.variable("cellsOnEdge", "i4", ("nCells", "TWO"))
    w.variable("odd", "i2", ("nVertLevels",))          # 6 bytes: exercises the 4-byte padding
    w.create()
    a = np.arange(21.0).reshape(7, 3) * 1.5
    b = np.arange(14, dtype=np.int32).reshape(7, 2)
    w.var("layerThickness")[:] = a
    w.var("cellsOnEdge")[:] = b
    w.var("odd")[:] = np.array([1, -2, 3])
    with netcdf_file(p, "r", mmap=False) as f:
        assert f.version_byte == 2
        assert np.array_equal(f.variables["layerThickness"][:], a)
        assert np.array_equal(f.variables["cellsOnEdge"][:], b)
        assert np.array_equal(f.variables["odd"][:], [1, -2, 3])
        assert f.on_a_sphere == b"YES" and float(f.sphere_radius) == 6371220.0
        assert f.variables["layerThickness"].units == b"m"
    # a file scipy wrote (the mesh file of meshgen.write_netcdf, the reference's input schema)
    m = meshgen.planar_hex(4, 4, 1000.0)
    q = str(tmp_path / "mesh.nc")
    meshgen.write_netcdf(m, q, {"layerThickness": np.full((m.nCells, 2), 5.0)})
    r = ncio.CDFReader(q)
    assert r.version == 2 and r.dim("nCells") == m.nCells
    assert np.array_equal(r.read("cellsOnEdge"), m.cellsOnEdge + 1)
    assert np.array_equal(r.read("weightsOnEdge"), m.weightsOnEdge)
    assert np.array_equal(r.read("layerThickness"), np.full((m.nCells, 2), 5.0))


@pytest.mark.parametrize("version", [1, 2, 5])
def test_round_trip_all_versions(tmp_path, version):
    p = str(tmp_path / "b.nc")
    w = ncio.CDFWriter(p, version)
    w.dim("x", 5); w.dim("y", 4)
    w.att("title", "round trip")
    w.variable("u", "f8", ("x", "y")); w.variable("i", "i4", ("y",)); w.variable("s", "f8", ())
    w.create()
    u = np.random.default_rng(1).standard_normal((5, 4))
    w.var("u")[:] = u; w.var("i")[:] = [4, 3, 2, 1]; w.var("s")[:] = 2.5
    r = ncio.CDFReader(p)
    assert r.version == version and r.atts["title"] == "round trip"
    assert np.array_equal(r.read("u"), u) and np.array_equal(r.read("i"), [4, 3, 2, 1]) and r.read("s")[0] == 2.5


def test_cdf5_holds_a_variable_beyond_4_gib_and_cdf2_refuses(tmp_path):
    """config-4-sized variable (sparse file: only the rows written take disk space)"""
    st = os.statvfs(str(tmp_path))
    if st.f_bavail * st.f_frsize < 1 << 30:
        pytest.skip("not enough scratch space")
    n, k = 2621442 * 3, 80                     # the edges x levels of BASELINE config 4: 5.03 GB
    for version in (2, 5):
        w = ncio.CDFWriter(str(tmp_path / f"big{version}.nc"), version)
        w.dim("NEdges", n); w.dim("NVertLayers", k); w.dim("NCells", 10)
        w.variable("NormalVelocity", "f8", ("NEdges", "NVertLayers"))
        w.variable("LayerThickness", "f8", ("NCells", "NVertLayers"))   # begins beyond 4 GiB
        if version == 2:
            with pytest.raises(ValueError):
                w.create()
            continue
        w.create()
        assert os.path.getsize(w.path) > 5 * 10**9
        if os.stat(w.path).st_blocks * 512 > 1 << 28:
            pytest.skip("file system does not keep the file sparse")
        rows = np.array([0, n // 2, n - 1])
        vals = np.arange(3 * k, dtype=float).reshape(3, k)
        w.var("NormalVelocity")[rows] = vals
        w.var("LayerThickness")[:] = 7.0
        r = ncio.CDFReader(w.path)
        assert r.vars["LayerThickness"]["begin"] > 1 << 32
        assert np.array_equal(np.asarray(r.var("NormalVelocity")[rows]), vals)
        assert np.all(np.asarray(r.var("LayerThickness")) == 7.0)
        os.remove(w.path)


def test_restart_two_ranks_scatter_then_resume_is_bit_identical(tmp_path):
    """Run 4 RK4 steps on one rank; separately run 2 steps, write a restart from TWO ranks' owned rows, read it
    back on one rank, run 2 more: the states are bit-identical and SimulationTime survives."""
    K, T = 3, 2
    m = cases.sphere_mesh(3)
    L1, cfg, (h, u, tr), dt = cases.sphere_case(3, K, T)
    co = lambda s: capi.rk_coeffs(capi.RK4, dt, t0=s * dt)
    ref = Oracle(L1, cfg, T); ref.make_state(h, u, tr)
    for s in range(4):
        ref.do_step(capi.RK4, co(s))
    want = [np.array(a) for a in ref.current()]
    ref.close()
    o = Oracle(L1, cfg, T); o.make_state(h, u, tr)
    for s in range(2):
        o.do_step(capi.RK4, co(s))
    gh, gu, gt = (np.array(a) for a in o.current())
    o.close()
    # global-order fields -> each of two ranks owns a part and scatters it by global ID
    cell_of, edge_of = L1.CellID[:-1] - 1, L1.EdgeID[:-1] - 1
    p = str(tmp_path / "restart.nc")
    ncio.create_restart(p, m.nCells, m.nEdges, K, T, sim_seconds=2 * dt)
    locs = decomp.build_all(m, 2, K, 3)
    inv_c = np.empty(m.nCells, np.int64); inv_c[cell_of] = np.arange(m.nCells)
    inv_e = np.empty(m.nEdges, np.int64); inv_e[edge_of] = np.arange(m.nEdges)
    for L in locs:
        ci, ei = L.CellID[:-1] - 1, L.EdgeID[:-1] - 1
        lh = np.zeros((L.NCellsSize, K)); lh[:-1] = gh[inv_c[ci]]
        lu = np.zeros((L.NEdgesSize, K)); lu[:-1] = gu[inv_e[ei]]
        lt = np.zeros((T, L.NCellsSize, K)); lt[:, :-1] = gt[:, inv_c[ci]]
        ncio.scatter_owned(p, L, lh, lu, lt)
    rh, ru, rt, stamp, secs = ncio.read_restart(p, L1, T)
    assert stamp == ncio.sim_time_string(2 * dt) and secs == 2 * dt
    assert np.array_equal(rh, gh) and np.array_equal(ru, gu) and np.array_equal(rt, gt)
    o = Oracle(L1, cfg, T); o.make_state(rh, ru, rt)
    for s in range(2, 4):
        o.do_step(capi.RK4, co(s))
    # the restarted run starts at CurTimeIndex 0; steps 2 and 3 of the uninterrupted run ran at time indices 0, 1 too
    for a, b in zip(o.current(), want):
        assert np.array_equal(a, b)
    o.close()


def test_sim_time_string_no_leap_calendar():
    assert ncio.sim_time_string(0) == "0001-01-01_00:00:00"
    assert ncio.sim_time_string(59 * 86400 + 3661) == "0001-03-01_01:01:01"      # no 29 February
    assert ncio.sim_time_string(365 * 86400) == "0002-01-01_00:00:00"


@pytest.mark.gpu
def test_restart_from_device_state_resumes_bit_identically(tmp_path):
    """checkpoint from DEVICE arrays: omega_field_to_host -> CDF-5 restart -> omega_field_to_device, fused tiled path"""
    from omega_b200 import model
    K, T = 16, 3
    m = cases.sphere_mesh(4)
    L, cfg, (h, u, tr), dt = cases.sphere_case(4, K, T)
    co = capi.rk_coeffs(capi.RK4, dt)
    mesh = model.HorzMesh(L)
    st = model.Stepper(mesh, cfg, T)
    a, b = model.OceanState(mesh, T), model.OceanState(mesh, T)
    a.copyToDevice(h, u, tr)
    for _ in range(4):
        st.doStep(a, capi.RK4, co, fused=True, compat=True)
    want = a.copyToHost(0)
    b.copyToDevice(h, u, tr)
    for _ in range(2):
        st.doStep(b, capi.RK4, co, fused=True, compat=True)
    p = str(tmp_path / "restart.nc")
    ncio.create_restart(p, m.nCells, m.nEdges, K, T, sim_seconds=2 * dt)
    ncio.scatter_owned(p, L, *b.copyToHost(0))
    rh, ru, rt, stamp, _ = ncio.read_restart(p, L, T)
    c = model.OceanState(mesh, T)
    c.copyToDevice(rh, ru, rt)
    for _ in range(2):
        st.doStep(c, capi.RK4, co, fused=True, compat=True)
    for x, y in zip(c.copyToHost(0), want):
        assert np.array_equal(x, y)
    st.close(); mesh.close()


@pytest.mark.parametrize("kind", ["land", "hept"])
def test_mesh_files_with_coastlines_and_heptagons(tmp_path, kind):
    """the reference's mesh file schema for the meshes with holes in their connectivity: missing neighbours are 0 in
    the file (1-based IDs), the rows of edgesOnEdge hold nEdgesOnEdge valid entries first; read back (scipy and the
    library-free reader) and decomposed, the local meshes equal those of the mesh in memory"""
    m = cases.land_mesh(3) if kind == "land" else cases.hept_mesh(3)
    p = str(tmp_path / "mesh.nc")
    meshgen.write_netcdf(m, p)
    r = ncio.CDFReader(p)
    assert r.dim("maxEdges") == m.maxEdges and r.dim("maxEdges2") == m.maxEdges2
    coe = r.read("cellsOnEdge")
    assert coe.min() == (0 if kind == "land" else 1) and np.array_equal(coe, m.cellsOnEdge + 1)
    assert np.array_equal(r.read("nEdgesOnEdge"), m.nEdgesOnEdge)
    back, _ = meshgen.read_netcdf(p)
    for k in ("cellsOnCell", "edgesOnCell", "verticesOnCell", "cellsOnEdge", "verticesOnEdge", "edgesOnEdge", "cellsOnVertex",
              "edgesOnVertex", "nEdgesOnCell", "nEdgesOnEdge", "weightsOnEdge", "kiteAreasOnVertex", "angleEdge", "bottomDepth"):
        assert np.array_equal(back[k], m[k]), k
    a, b = decomp.build_all(m, 3, 2)[1], decomp.build_all(back, 3, 2)[1]
    for k in ("CellID", "EdgeID", "VertexID", "CellsOnEdge", "EdgesOnEdge", "CellsOnVertex", "WeightsOnEdge", "AreaCell"):
        assert np.array_equal(a[k], b[k]), k
