"""The C++ host mirror (omega_b200/host/OmegaB200.h: HorzMesh, OceanState,
Tracers, AuxiliaryState, Tendencies, TimeStepper with the reference's method
names) driven by a small C++ program, checked against the oracle."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from omega_b200 import capi
from tests import cases
from tests.oracle_lib import Oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "host_mirror_test")


def build_exe():
    src = os.path.join(ROOT, "tests", "cpp", "host_mirror_test.cpp")
    lib_dir = os.path.join(ROOT, "omega_b200", "csrc")
    cmd = ["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "omega_b200", "host"),
           "-I/usr/local/cuda/include", src, "-o", EXE, "-L" + lib_dir, "-lomega_b200", "-L/usr/local/cuda/lib64", "-lcudart", "-lcublas",
           "-Wl,-rpath," + lib_dir, "-Wl,-rpath,/usr/local/cuda/lib64"]
    subprocess.run(cmd, check=True)


def test_host_mirror_compiles_against_the_c_abi():
    """CPU check: the mirror header and the test driver compile and link."""
    if not os.path.exists(capi.LIB_PATH):
        pytest.skip("libomega_b200.so not built")
    build_exe()
    assert os.path.exists(EXE)


def write_case(d, L, cfg, h, u, tr, dt, stepper, fused, nsteps):
    """mesh, config and state of one rank as raw files for tests/cpp/host_mirror_test"""
    T = tr.shape[0]
    os.makedirs(d, exist_ok=True)
    for name, _ in capi.MeshDesc._fields_[13:]:
        np.ascontiguousarray(L[name]).astype(np.int32 if name in (
            "NEdgesOnCell", "EdgesOnCell", "CellsOnEdge", "VerticesOnEdge", "NEdgesOnEdge", "EdgesOnEdge",
            "CellsOnVertex", "EdgesOnVertex") else np.float64).tofile(os.path.join(d, name + ".bin"))
    with open(os.path.join(d, "meta.txt"), "w") as f:
        for name, _ in capi.MeshDesc._fields_[:13]:
            f.write(f"{name} {int(L[name])}\n")
        f.write(f"NTracers {T}\nNSteps {nsteps}\nDt {dt}\nStepper {stepper}\nFused {fused}\n")
    with open(os.path.join(d, "cfg.bin"), "wb") as f:
        f.write(bytes(cfg))
    h.tofile(os.path.join(d, "h.bin")); u.tofile(os.path.join(d, "u.bin")); tr.tofile(os.path.join(d, "tr.bin"))


@pytest.mark.gpu
def test_host_mirror_two_ranks_step_without_python_transport(tmp_path):
    """Two C++ host processes, one GPU each, step RK4 through TimeStepper::doStep: the exchange rounds run
    inside omega_stepper_do_step over NCCL (no Python, no caller-supplied transport).  Every local
    element of both ranks against the 1-rank oracle."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    from omega_b200 import analytic, decomp, parallel
    build_exe()
    K, T, nsteps, world = 16, 2, 3, 2
    mesh_g = cases.sphere_mesh(4)
    L1, cfg, (h, u, tr), dt = cases.sphere_case(4, K, T)
    locs = decomp.build_all(mesh_g, world, K, 4)
    lists = parallel.build_halo_lists(locs)
    for r, (L, hl) in enumerate(zip(locs, lists)):
        d = str(tmp_path / f"r{r}")
        lh, lu, ltr = analytic.tc5(L, T)
        write_case(d, L, cfg, lh, lu, ltr, dt, capi.RK4, 1, nsteps)
        np.array(hl["neighbors"], np.int32).tofile(os.path.join(d, "halo_neighbors.bin"))
        flat = parallel.flatten_lists(hl)
        for dirn, tag in (("send", "send"), ("recv", "recv")):
            for kind in parallel.KINDS:
                off, idx = flat[(dirn, kind)]
                off.tofile(os.path.join(d, f"halo_{tag}_off_{kind}.bin"))
                idx.tofile(os.path.join(d, f"halo_{tag}_idx_{kind}.bin"))
    procs = [subprocess.Popen([EXE, str(tmp_path / f"r{r}"), str(r), str(world)]) for r in range(world)]
    for p in procs:
        assert p.wait(timeout=300) == 0
    o = Oracle(L1, cfg, T)
    o.make_state(h, u, tr)
    for s in range(nsteps):
        o.do_step(capi.RK4, capi.rk_coeffs(capi.RK4, dt, t0=s * dt))
    gh, gu, gt = o.current()
    cell_of, edge_of = L1.CellID[:-1] - 1, L1.EdgeID[:-1] - 1
    H = np.empty_like(gh[:-1]); H[cell_of] = gh[:-1]
    U = np.empty_like(gu[:-1]); U[edge_of] = gu[:-1]
    TR = np.empty_like(gt[:, :-1]); TR[:, cell_of] = gt[:, :-1]
    for r, L in enumerate(locs):
        d = str(tmp_path / f"r{r}")
        lh = np.fromfile(os.path.join(d, "out_h.bin")).reshape(-1, K)
        lu = np.fromfile(os.path.join(d, "out_u.bin")).reshape(-1, K)
        lt = np.fromfile(os.path.join(d, "out_tr.bin")).reshape(T, -1, K)
        ci, ei = L.CellID[:L.NCellsAll] - 1, L.EdgeID[:L.NEdgesAll] - 1
        for a, b in ((lh[:L.NCellsAll], H[ci]), (lu[:L.NEdgesAll], U[ei]), (lt[:, :L.NCellsAll], TR[:, ci])):
            assert cases.rel_l2(a, b) <= 1e-10, (r, cases.rel_l2(a, b))
    o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("stepper,fused,custom,wind", [(capi.RK4, 1, 0, 0), (capi.RK4, 0, 0, 0), (capi.FB, 0, 0, 0), (capi.RK4, 1, 1, 0),
                                                       (capi.RK4, 1, 0, 1)])
def test_host_mirror_matches_oracle(tmp_path, stepper, fused, custom, wind):
    """custom = 1: a std::function custom velocity tendency installed on the Tendencies mirror as the reference's
    TimeStepperTest installs its own (Tendencies.h:51-53), on top of every built-in term.  wind = 1: wind forcing,
    bottom drag and upwind fluxes on, the stress arrays through AuxiliaryState::setWindStress / exchangeHalo"""
    build_exe()
    if wind:
        L, cfg, (h, u, tr), dt, (zonal, merid) = cases.all_terms_case(3, 6, 2)
        zonal.tofile(os.path.join(str(tmp_path), "zonal.bin")); merid.tofile(os.path.join(str(tmp_path), "merid.bin"))
    else:
        L, cfg, (h, u, tr), dt = cases.sphere_case(3, 6, 2)
    T, nsteps = tr.shape[0], 3
    desc, keep = capi.make_mesh_desc(L)
    d = str(tmp_path)
    for name, _ in capi.MeshDesc._fields_[13:]:
        np.ascontiguousarray(L[name]).astype(np.int32 if name in (
            "NEdgesOnCell", "EdgesOnCell", "CellsOnEdge", "VerticesOnEdge", "NEdgesOnEdge", "EdgesOnEdge",
            "CellsOnVertex", "EdgesOnVertex") else np.float64).tofile(os.path.join(d, name + ".bin"))
    with open(os.path.join(d, "meta.txt"), "w") as f:
        for name, _ in capi.MeshDesc._fields_[:13]:
            f.write(f"{name} {int(L[name])}\n")
        f.write(f"NTracers {T}\nNSteps {nsteps}\nDt {dt}\nStepper {stepper}\nFused {fused}\nCustom {custom}\nWind {wind}\n")
    with open(os.path.join(d, "cfg.bin"), "wb") as f:
        f.write(bytes(cfg))
    h.tofile(os.path.join(d, "h.bin")); u.tofile(os.path.join(d, "u.bin")); tr.tofile(os.path.join(d, "tr.bin"))
    subprocess.run([EXE, d], check=True)
    o = Oracle(L, cfg, T)
    if wind:
        o.set_wind_stress(zonal, merid)
    if custom:
        def decay(tend, hh, uu, time):
            tend -= 1.0e-5 * uu
        o.set_custom_tendencies(None, decay)
    o.tend_all(h, u, tr)
    for nm, key in (("out_htend", "LayerThicknessTend"), ("out_utend", "NormalVelocityTend"), ("out_trtend", "TracerTend")):
        got = np.fromfile(os.path.join(d, nm + ".bin")).reshape(o.tend[key].shape)
        assert cases.rel_max(got, o.tend[key]) <= 1e-12, nm
    o.make_state(h, u, tr)
    for s in range(nsteps):
        o.do_step(stepper, capi.rk_coeffs(stepper, dt, t0=s * dt))
    for nm, exp in zip(("out_h", "out_u", "out_tr"), o.current()):
        got = np.fromfile(os.path.join(d, nm + ".bin")).reshape(exp.shape)
        assert cases.rel_l2(got, exp) <= 1e-10, nm
    o.close()
    _check_vertical_mirror(d, L, np.fromfile(os.path.join(d, "out_h.bin")).reshape(h.shape))


def _check_vertical_mirror(d, L, hnow):
    """OmegaB200Vertical.h (VertCoord, Eos, TriDiagSolver, TriDiagDiffSolver) in the C++ driver against the C oracle,
    on the inputs the driver builds from the stepped thickness (tests/cpp/host_mirror_test.cpp)."""
    import ctypes as C
    from tests import oracle_lib
    if not os.path.exists(oracle_lib.STRICT):
        oracle_lib.build_oracle()
    lib = C.CDLL(oracle_lib.STRICT)
    P = lambda a: a.ctypes.data_as(capi.c_dbl_p)
    PI = lambda a: a.ctypes.data_as(capi.c_int_p)
    n, K = hnow.shape
    i = np.arange(n)
    lo = np.where((i % 3 == 0) & (K > 3), 1, 0).astype(np.int32)
    hi = np.where((i % 5 == 0) & (K > 3), K - 2, K - 1).astype(np.int32)
    k = np.arange(K)[None, :]
    ct = 2.0 + 0.01 * (i % 100)[:, None] + 0.5 * k
    sa = 33.0 + 0.02 * (i % 50)[:, None] + 0.1 * k
    surf, tid, sal = 100.0 + i % 7, 0.01 * (i % 11), 0.002 * (i % 13)
    ref_h, wt = np.ascontiguousarray(hnow * 0.9), 1.0 + 0.1 * np.arange(K)
    rho0 = C.c_double(1026.0)
    bot = np.ascontiguousarray(L.BottomDepth, np.float64)
    pint, pmid = np.zeros((n, K + 1)), np.zeros((n, K))
    lib.oracle_vert_pressure(n, K, PI(lo), PI(hi), rho0, P(hnow), P(np.ascontiguousarray(surf, np.float64)), P(pint), P(pmid))
    pd = np.ascontiguousarray(pmid * 1.0e-4)
    sv, svd = np.zeros((n, K)), np.zeros((n, K))
    lib.oracle_eos_specvol(n, K, 1, 0, C.c_double(-0.2), C.c_double(0.8), C.c_double(1000.0), P(np.ascontiguousarray(ct)), P(np.ascontiguousarray(sa)), P(pd), P(sv))
    lib.oracle_eos_specvol(n, K, 1, 1, C.c_double(-0.2), C.c_double(0.8), C.c_double(1000.0), P(np.ascontiguousarray(ct)), P(np.ascontiguousarray(sa)), P(pd), P(svd))
    zint, zmid, geo, tgt = np.zeros((n, K + 1)), np.zeros((n, K)), np.zeros((n, K)), np.zeros((n, K))
    lib.oracle_vert_zheight(n, K, PI(lo), PI(hi), rho0, P(bot), P(hnow), P(sv), P(zint), P(zmid))
    lib.oracle_vert_geopotential(n, K, PI(lo), PI(hi), P(zmid), P(np.ascontiguousarray(tid, np.float64)), P(np.ascontiguousarray(sal, np.float64)), P(geo))
    lib.oracle_vert_target(n, K, PI(lo), PI(hi), rho0, P(pint), P(ref_h), P(np.ascontiguousarray(wt)), P(tgt))
    nall = int(L.NCellsAll)
    active = (k >= lo[:, None]) & (k <= hi[:, None])
    active[nall:] = False
    act1 = np.zeros((n, K + 1), bool)
    act1[:, :-1] |= active
    act1[:, 1:] |= active
    for nm, exp, m in (("out_pint", pint, act1), ("out_pmid", pmid, active), ("out_specvol", sv, None), ("out_specvol_disp", svd, None),
                       ("out_zint", zint, act1), ("out_zmid", zmid, active), ("out_geop", geo, active), ("out_target", tgt, active)):
        got = np.fromfile(os.path.join(d, nm + ".bin")).reshape(exp.shape)
        if m is None:
            m = np.zeros(exp.shape, bool)
            m[:nall] = True
        err = np.abs(got - exp)[m].max() / np.abs(exp[m]).max()
        assert err <= 1e-12, (nm, err)
    # tridiagonal solvers: the driver's systems, solved by the oracle's Thomas forms
    g = np.fromfile(os.path.join(d, "in_tridiag_g.bin")).reshape(nall, K)
    hd = np.fromfile(os.path.join(d, "in_tridiag_h.bin")).reshape(nall, K)
    x = np.fromfile(os.path.join(d, "in_tridiag_x.bin")).reshape(nall, K)
    xr, hh, al = x.copy(), hd.copy(), np.zeros((nall, K))
    lib.oracle_tridiag_diffusion_thomas(nall, K, P(g), P(hh), P(xr), P(al))
    for nm in ("out_tridiag_diff", "out_tridiag"):
        got = np.fromfile(os.path.join(d, nm + ".bin")).reshape(nall, K)
        assert np.abs(got - xr).max() / np.abs(xr).max() <= 1e-12, nm
