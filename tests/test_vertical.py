"""the Omega-1 column pieces -- VertCoord integrals, equation of state, tridiagonal solvers.
CPU: the C oracle (oracle/vertical_oracle.c) against the reference's own Eos functors (oracle/_ref), the published
check values of test/ocn/EosTest.cpp:39-47, closed forms and dense solves.  GPU: csrc/vertical.cu against the oracle."""
import ctypes as C
import os

import numpy as np
import pytest

from omega_b200 import capi
from tests import cases, oracle_lib

dp, ip = capi.c_dbl_p, capi.c_int_p
P = lambda a: a.ctypes.data_as(dp)
PI = lambda a: a.ctypes.data_as(ip)


def oracle():
    if not os.path.exists(oracle_lib.STRICT):
        oracle_lib.build_oracle()
    lib = C.CDLL(oracle_lib.STRICT)
    lib.oracle_teos10_specvol.restype = C.c_double
    return lib


def column_case(n=37, K=60, seed=3):
    rng = np.random.default_rng(seed)
    kmin = rng.integers(0, 3, n).astype(np.int32)
    kmax = rng.integers(K - 20, K, n).astype(np.int32)
    kmin[0], kmax[0] = 0, K - 1
    h = 5.0 + 40.0 * rng.random((n, K))
    surf = 1.0e4 * rng.random(n)
    bot = 500.0 + 4000.0 * rng.random(n)
    ct = -2.0 + 30.0 * rng.random((n, K))
    sa = 30.0 + 8.0 * rng.random((n, K))
    return kmin, kmax, h, surf, bot, ct, sa


def test_eos_published_values_and_reference_functor_bit_exact():
    lib = oracle()
    d = C.c_double
    v = lib.oracle_teos10_specvol(d(30.0), d(10.0), d(1000.0))
    assert abs(v - 0.0009732819628) <= 1e-10 * 0.0009732819628          # EosTest.cpp:40,47 (RTol 1e-10)
    ct, sa, p, out = (np.full((1, 1), x) for x in (10.0, 30.0, 1000.0, 0.0))
    lib.oracle_eos_specvol(1, 1, 0, 0, d(-0.2), d(0.8), d(1000.0), P(ct), P(sa), P(p), P(out))
    assert abs(out[0, 0] - 0.0009784735812133072) <= 1e-10 * 0.00097847       # EosTest.cpp:41-42
    if not os.path.exists(oracle_lib.REF):
        pytest.skip("oracle/_ref not built")
    ref = C.CDLL(oracle_lib.REF)
    kmin, kmax, h, surf, bot, ct, sa = column_case()
    n, K = ct.shape
    pr = np.cumsum(h, 1) * 1.0    # dbar-like, increasing with depth
    for eos, kdisp in ((1, 0), (1, 1), (1, -2), (0, 0)):
        a, b = np.zeros((n, K)), np.zeros((n, K))
        lib.oracle_eos_specvol(n, K, eos, kdisp, d(-0.2), d(0.8), d(1000.0), P(ct), P(sa), P(pr), P(a))
        ref.ref_eos_specvol(n, K, eos, kdisp, P(ct), P(sa), P(pr), P(b))
        assert np.array_equal(a, b), (eos, kdisp, np.abs(a - b).max())


def test_vert_coord_oracle_closed_forms():
    """uniform columns, as test/ocn/VertCoordTest.cpp checks them: p = ps + g rho0 sum h, z from the bottom up"""
    lib = oracle()
    n, K, rho0, g = 5, 12, 1026.0, 9.80616
    kmin, kmax = np.zeros(n, np.int32), np.full(n, K - 1, np.int32)
    kmin[2], kmax[3] = 2, K - 4
    h = np.full((n, K), 10.0)
    surf, bot = np.full(n, 300.0), np.full(n, 1000.0)
    pint, pmid = np.zeros((n, K + 1)), np.zeros((n, K))
    lib.oracle_vert_pressure(n, K, PI(kmin), PI(kmax), C.c_double(rho0), P(h), P(surf), P(pint), P(pmid))
    for i in range(n):
        for k in range(kmin[i], kmax[i] + 1):
            assert abs(pint[i, k + 1] - (300.0 + g * rho0 * 10.0 * (k + 1 - kmin[i]))) < 1e-8
            assert abs(pmid[i, k] - 0.5 * (pint[i, k] + pint[i, k + 1])) < 1e-8
    sv = np.full((n, K), 1.0 / rho0)
    zint, zmid = np.zeros((n, K + 1)), np.zeros((n, K))
    lib.oracle_vert_zheight(n, K, PI(kmin), PI(kmax), C.c_double(rho0), P(bot), P(h), P(sv), P(zint), P(zmid))
    for i in range(n):
        assert zint[i, kmax[i] + 1] == -1000.0
        assert abs(zint[i, kmin[i]] - (-1000.0 + 10.0 * (kmax[i] - kmin[i] + 1))) < 1e-9
    geo, tid, sal = np.zeros((n, K)), np.full(n, 0.5), np.full(n, 0.25)
    lib.oracle_vert_geopotential(n, K, PI(kmin), PI(kmax), P(zmid), P(tid), P(sal), P(geo))
    assert abs(geo[0, 3] - (g * zmid[0, 3] + 0.75)) < 1e-12
    # target thickness: uniform weights and reference thickness equal to the column -> target == reference
    wgt, tgt = np.ones(K), np.zeros((n, K))
    lib.oracle_vert_target(n, K, PI(kmin), PI(kmax), C.c_double(rho0), P(pint), P(h), P(wgt), P(tgt))
    for i in range(n):
        assert np.allclose(tgt[i, kmin[i]:kmax[i] + 1], 10.0, rtol=1e-13)


def tridiag_case(nb=9, nr=60, seed=5):
    rng = np.random.default_rng(seed)
    dl, du = -rng.random((nb, nr)), -rng.random((nb, nr))
    d = 1.0 + rng.random((nb, nr)) - dl - du
    x = rng.standard_normal((nb, nr))
    return dl, d, du, x


def dense_solve(dl, d, du, x):
    out = np.zeros_like(x)
    for b in range(x.shape[0]):
        A = np.diag(d[b]) + np.diag(dl[b, 1:], -1) + np.diag(du[b, :-1], 1)
        out[b] = np.linalg.solve(A, x[b])
    return out


def test_tridiag_oracle_against_dense_solve_and_stability_scenario():
    lib = oracle()
    dl, d, du, x = tridiag_case()
    want = dense_solve(dl, d, du, x)
    dd, xx = d.copy(), x.copy()
    lib.oracle_tridiag_thomas(x.shape[0], x.shape[1], P(dl), P(dd), P(du), P(xx))
    assert np.abs(xx - want).max() <= 1e-12 * np.abs(want).max()
    # diffusion form == the general form built from the same G, H
    rng = np.random.default_rng(1)
    nb, nr = 4, 50
    g, h, rhs = rng.random((nb, nr)) * 5.0, 1.0 + rng.random((nb, nr)), rng.standard_normal((nb, nr))
    g[:, -1] = 0.0
    gm = np.concatenate([np.zeros((nb, 1)), g[:, :-1]], 1)
    want = dense_solve(-gm, h + gm + g, -g, rhs)
    hh, xx, al = h.copy(), rhs.copy(), np.zeros(nr)
    lib.oracle_tridiag_diffusion_thomas(nb, nr, P(g), P(hh), P(xx), P(al))
    assert np.abs(xx - want).max() <= 1e-12 * np.abs(want).max()
    # abrupt 1e14 jump in diffusivity (TriDiagSolversTest.cpp:447-470): the diffusion form stays finite
    g2 = g.copy(); g2[:, nr // 2:] = 1e14; g2[:, -1] = 0.0
    hh, xx = h.copy(), rhs.copy()
    lib.oracle_tridiag_diffusion_thomas(nb, nr, P(g2), P(hh), P(xx), P(al))
    assert np.isfinite(xx).all()


# ---------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
def test_gpu_vertical_pieces_against_oracle():
    import torch
    from omega_b200 import model, vertical
    lib_o = oracle()
    K = 60
    L, cfg, (h0, u0, tr0), dt = cases.sphere_case(3, K, 0)
    mesh = model.HorzMesh(L)
    n = int(L.NCellsAll)
    kmin, kmax, h, surf, bot, ct, sa = column_case(n + 1, K)
    kmin[-1], kmax[-1] = 0, 0
    botd = np.asarray(L.BottomDepth, np.float64)
    d = C.c_double
    rho0 = 1026.0
    # oracle (host order, NCellsAll columns)
    pint, pmid = np.zeros((n + 1, K + 1)), np.zeros((n + 1, K))
    lib_o.oracle_vert_pressure(n, K, PI(kmin), PI(kmax), d(rho0), P(h), P(surf), P(pint), P(pmid))
    sv = np.zeros((n + 1, K))
    lib_o.oracle_eos_specvol(n, K, 1, 0, d(-0.2), d(0.8), d(1000.0), P(ct), P(sa), P(np.ascontiguousarray(pmid * 1e-4)), P(sv))
    svd = np.zeros((n + 1, K))
    lib_o.oracle_eos_specvol(n, K, 1, 1, d(-0.2), d(0.8), d(1000.0), P(ct), P(sa), P(np.ascontiguousarray(pmid * 1e-4)), P(svd))
    svl = np.zeros((n + 1, K))
    lib_o.oracle_eos_specvol(n, K, 0, 0, d(-0.2), d(0.8), d(1000.0), P(ct), P(sa), P(pmid), P(svl))
    zint, zmid = np.zeros((n + 1, K + 1)), np.zeros((n + 1, K))
    lib_o.oracle_vert_zheight(n, K, PI(kmin), PI(kmax), d(rho0), P(botd), P(h), P(sv), P(zint), P(zmid))
    tid, sal = np.linspace(0, 1, n + 1), np.linspace(1, 2, n + 1)
    geo = np.zeros((n + 1, K))
    lib_o.oracle_vert_geopotential(n, K, PI(kmin), PI(kmax), P(zmid), P(tid), P(sal), P(geo))
    refh, wgt, tgt = 3.0 + h * 0.9, 0.5 + np.linspace(0, 1, K), np.zeros((n + 1, K))
    lib_o.oracle_vert_target(n, K, PI(kmin), PI(kmax), d(rho0), P(pint), P(refh), P(wgt), P(tgt))
    # GPU
    vc = vertical.VertCoord(mesh, kmin, kmax, rho0)
    dev = lambda a: mesh.to_device(capi.ON_CELL, np.ascontiguousarray(a))
    dh, dct, dsa = dev(h), dev(ct), dev(sa)
    vc.computePressure(dh, dev(surf))
    eos = vertical.Eos(mesh, vertical.TEOS10_EOS)
    dp_dbar = vc.PressureMid * 1e-4
    eos.computeSpecVol(dct, dsa, dp_dbar)
    eos.computeSpecVolDisp(dct, dsa, dp_dbar, 1)
    lin = vertical.Eos(mesh, vertical.LINEAR_EOS)
    lin.computeSpecVol(dct, dsa, None)
    vc.computeZHeight(dh, eos.SpecVol)
    vc.computeGeopotential(dev(tid), dev(sal))
    vc.computeTargetThickness(dev(refh), torch.from_numpy(wgt).to(mesh.device))
    host = lambda t: mesh.to_host(capi.ON_CELL, t)
    act = np.zeros((n + 1, K), bool)
    for i in range(n):
        act[i, kmin[i]:kmax[i] + 1] = True
    acti = np.zeros((n + 1, K + 1), bool)
    for i in range(n):
        acti[i, kmin[i]:kmax[i] + 2] = True

    def close(got, want, mask, what):
        err = np.abs(got[mask] - want[mask]).max() / np.abs(want[mask]).max()
        assert err <= 1e-12, (what, err)

    close(host(vc.PressureInterface), pint, acti, "PressureInterface")
    close(host(vc.PressureMid), pmid, act, "PressureMid")
    allm = np.ones((n + 1, K), bool); allm[-1] = False
    close(host(eos.SpecVol), sv, allm, "SpecVol TEOS-10")
    close(host(eos.SpecVolDisplaced), svd, allm, "SpecVolDisplaced TEOS-10")
    close(host(lin.SpecVol), svl, allm, "SpecVol linear")
    close(host(vc.ZInterface), zint, acti, "ZInterface")
    close(host(vc.ZMid), zmid, act, "ZMid")
    close(host(vc.GeopotentialMid), geo, act, "GeopotentialMid")
    close(host(vc.LayerThicknessTarget), tgt, act, "LayerThicknessTarget")
    # the published TEOS-10 check value on the device (EosTest.cpp:40-47)
    one = lambda v: mesh.to_device(capi.ON_CELL, np.full((n + 1, K), v))
    eos.computeSpecVol(one(10.0), one(30.0), one(1000.0))
    got = host(eos.SpecVol)[0, 0]
    assert abs(got - 0.0009732819628) <= 1e-10 * got
    mesh.close()


@pytest.mark.gpu
@pytest.mark.parametrize("nr", [1, 2, 33, 60, 64, 100])
def test_gpu_tridiagonal_solvers(nr):
    import torch
    from omega_b200 import vertical
    lib = capi.load_library()
    lib_o = oracle()
    dl, d, du, x = tridiag_case(nb=257, nr=nr)
    want = x.copy(); dd = d.copy()
    lib_o.oracle_tridiag_thomas(x.shape[0], nr, P(dl), P(dd), P(du), P(want))
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).cuda()
    xg = t(x)
    vertical.tridiag_solve(lib, t(dl), t(d), t(du), xg)
    assert np.abs(xg.cpu().numpy() - want).max() <= 1e-12 * np.abs(want).max()
    rng = np.random.default_rng(2)
    g, h, rhs = rng.random((257, nr)) * 5.0, 1.0 + rng.random((257, nr)), rng.standard_normal((257, nr))
    g[:, -1] = 0.0
    for jump in (False, True):
        gg = g.copy()
        if jump and nr > 2:
            gg[:, nr // 2:] = 1e14; gg[:, -1] = 0.0
        hh, want, al = h.copy(), rhs.copy(), np.zeros(nr)
        lib_o.oracle_tridiag_diffusion_thomas(257, nr, P(gg), P(hh), P(want), P(al))
        xg = t(rhs)
        vertical.tridiag_diffusion_solve(lib, t(gg), t(h), xg)
        got = xg.cpu().numpy()
        assert np.isfinite(got).all()
        assert np.abs(got - want).max() <= 1e-11 * np.abs(want).max(), (nr, jump, np.abs(got - want).max())


def test_vert_coord_oracle_equals_the_reference_vertcoord():
    """oracle/vertical_oracle.c against the reference's own src/ocn/VertCoord.cpp (compiled unmodified,
    oracle/_ref/libomega_ref_vertcoord.so: Kokkos team scans as one-thread teams in index order): pressure, z height,
    geopotential and target thickness of random columns with partial layer ranges, bit for bit."""
    ref_path = os.path.join(os.path.dirname(oracle_lib.REF), "libomega_ref_vertcoord.so")
    if not os.path.exists(ref_path):
        pytest.skip("oracle/_ref/libomega_ref_vertcoord.so not built")
    lib, ref = oracle(), C.CDLL(ref_path)
    kmin, kmax, h, surf, bot, ct, sa = column_case(n=53, K=40, seed=11)
    n, K = h.shape
    rng = np.random.default_rng(2)
    sv = 9.6e-4 + 2.0e-5 * rng.random((n, K))
    tid, sal = 0.3 * rng.random(n), 0.05 * rng.random(n)
    refh, wt = np.ascontiguousarray(h * (0.8 + 0.4 * rng.random((n, K)))), 0.5 + rng.random(K)
    d = C.c_double(1026.0)
    o = {k: np.zeros((n, K + 1)) for k in ("pint", "zint")}
    o.update({k: np.zeros((n, K)) for k in ("pmid", "zmid", "geo", "tgt")})
    r = {k: np.zeros_like(v) for k, v in o.items()}
    lib.oracle_vert_pressure(n, K, PI(kmin), PI(kmax), d, P(h), P(surf), P(o["pint"]), P(o["pmid"]))
    lib.oracle_vert_zheight(n, K, PI(kmin), PI(kmax), d, P(bot), P(h), P(sv), P(o["zint"]), P(o["zmid"]))
    lib.oracle_vert_geopotential(n, K, PI(kmin), PI(kmax), P(o["zmid"]), P(tid), P(sal), P(o["geo"]))
    lib.oracle_vert_target(n, K, PI(kmin), PI(kmax), d, P(o["pint"]), P(refh), P(wt), P(o["tgt"]))
    assert ref.ref_vertcoord_run(n, K, PI(kmin), PI(kmax), d, P(bot), P(h), P(surf), P(sv), P(tid), P(sal), P(refh), P(wt),
                                 P(r["pint"]), P(r["pmid"]), P(r["zint"]), P(r["zmid"]), P(r["geo"]), P(r["tgt"])) == 0
    k = np.arange(K)[None, :]
    act = (k >= kmin[:, None]) & (k <= kmax[:, None])
    act1 = np.zeros((n, K + 1), bool)
    act1[:, :-1] |= act
    act1[:, 1:] |= act
    for nm in o:
        m = act1 if o[nm].shape[1] == K + 1 else act
        assert np.abs(r[nm][m]).max() > 0
        assert np.array_equal(o[nm][m], r[nm][m]), (nm, np.abs(o[nm][m] - r[nm][m]).max())


def test_tridiagonal_oracle_equals_the_reference_thomas_solvers():
    """oracle_tridiag_thomas / oracle_tridiag_diffusion_thomas against the reference's own ThomasSolver::solve and
    ThomasDiffusionSolver::solve (src/base/TriDiagSolvers.h, compiled unmodified), bit for bit -- including the system
    with a 1e14 jump in diffusivity of test/base/TriDiagSolversTest.cpp:428-470."""
    ref_path = os.path.join(os.path.dirname(oracle_lib.REF), "libomega_ref_vertcoord.so")
    if not os.path.exists(ref_path):
        pytest.skip("oracle/_ref/libomega_ref_vertcoord.so not built")
    lib, ref = oracle(), C.CDLL(ref_path)
    dl, d, du, x = tridiag_case(nb=7, nr=50, seed=9)
    xo, do = x.copy(), d.copy()
    lib.oracle_tridiag_thomas(7, 50, P(dl), P(do), P(du), P(xo))
    xr = x.copy()
    assert ref.ref_tridiag_thomas(7, 50, P(dl), P(d), P(du), P(xr)) == 0
    assert np.array_equal(xo, xr), np.abs(xo - xr).max()
    rng = np.random.default_rng(4)
    g, h, x2 = rng.random((5, 40)) * 3.0, 0.5 + rng.random((5, 40)), rng.standard_normal((5, 40))
    g[2, 17] = 1.0e14                      # abrupt jump in diffusivity
    g[:, -1] = 0.0
    xo, ho, al = x2.copy(), h.copy(), np.zeros((5, 40))
    lib.oracle_tridiag_diffusion_thomas(5, 40, P(g), P(ho), P(xo), P(al))
    xr = x2.copy()
    assert ref.ref_tridiag_diffusion_thomas(5, 40, P(g), P(h), P(xr)) == 0
    assert np.all(np.isfinite(xr)) and np.array_equal(xo, xr), np.abs(xo - xr).max()
