"""Invariants of the algorithm on the path (doc/design/OmegaV0ShallowWater.md:61-67: volume and tracer content to
round-off, a tracer == 1 stays 1; the TRiSK operators conserve total energy in space, so an inviscid run loses energy
only through the time stepping, RK4: fifth order).  Run on the C oracle, i.e. on the reference's arithmetic; on a
closed sphere and on the mesh with coastlines (no normal flow through the coast: u = 0 on the masked edges, which the
masked tendencies keep).  The discrete energy the operators conserve:
   E = sum_e dc dv / 2 * hbar_e u_e^2  +  sum_c A_c g (h_c - BottomDepth_c)^2 / 2,   hbar_e = (h_c0 + h_c1) / 2
(KineticAuxVars.h:20-47 defines KE_c = sum_e dc dv u^2 / 4 / A_c, whose thickness-weighted sum over cells is the first
term; SSHGradOnEdge works on h - BottomDepth).  Time levels mapped through the ring (RawTimeLevelCompat = 0): with the
reference's raw indexing the tracer update is not consistent with the thickness (tests/test_tc2_steady.py)."""
import numpy as np
import pytest

from omega_b200 import capi
from tests import cases
from tests.oracle_lib import Oracle

GRAV = 9.80665
INVISCID = dict(VelDiffTendencyEnable=0, VelHyperDiffTendencyEnable=0, TracerDiffTendencyEnable=0, TracerHyperDiffTendencyEnable=0)


def energy(L, h, u):
    n, ne = L.NCellsOwned, L.NEdgesOwned
    hbar = 0.5 * (h[L.CellsOnEdge[:ne, 0]] + h[L.CellsOnEdge[:ne, 1]])
    ke = ((L.DcEdge[:ne] * L.DvEdge[:ne] / 2)[:, None] * hbar * u[:ne] ** 2).sum(0)
    eta = h[:n] - L.BottomDepth[:n, None]
    return ke + (L.AreaCell[:n, None] * GRAV * eta ** 2 / 2).sum(0), ke


def run(case, fac, hours=6.0, T=2):
    L, cfg, (h, u, tr), dt = case(3, 2, T, **INVISCID)
    coast = L.CellsOnEdge[:, 1] == L.NCellsAll
    u = u.copy(); u[coast] = 0.0
    dt *= fac
    o = Oracle(L, cfg, T)
    o.make_state(h, u, tr)
    n, A = L.NCellsOwned, L.AreaCell[:L.NCellsOwned, None]
    e0, k0 = energy(L, h, u)
    v0, q0 = (A * h[:n]).sum(0), (A * h[:n] * tr[0, :n]).sum(0)
    for s in range(int(hours * 3600 / dt)):
        o.do_step(capi.RK4, capi.rk_coeffs(capi.RK4, dt, t0=s * dt), compat=0)
    gh, gu, gt = o.current()
    e1, k1 = energy(L, gh, gu)
    out = dict(energy=float(np.abs((e1 - e0) / e0).max()), exchange=float(np.abs((k1 - k0) / e0).max()),
               volume=float(np.abs(((A * gh[:n]).sum(0) - v0) / v0).max()),
               tracer=float(np.abs(((A * gh[:n] * gt[0, :n]).sum(0) - q0) / q0).max()),
               one=float(np.abs(gt[-1, :n] - 1.0).max()), coast=float(np.abs(gu[coast]).max()) if coast.any() else 0.0)
    o.close()
    return out


@pytest.mark.parametrize("case", [cases.sphere_case, cases.land_case], ids=["closed_sphere", "coastline"])
def test_volume_tracer_and_energy(case):
    a, b = run(case, 1.0), run(case, 0.5)
    print(case.__name__, a, b)
    for r in (a, b):
        assert r["volume"] <= 5e-15 and r["tracer"] <= 5e-15      # round-off (OmegaV0ShallowWater.md:61-65)
        assert r["one"] == 0.0                                    # tracer == 1 stays 1 to the bit (:66-67)
        assert r["coast"] == 0.0                                  # masked edges keep their velocity
        assert r["exchange"] > 1e-3                               # kinetic and potential energy do trade
    # spatially exact: what is lost is the time stepping's, and it falls as dt^5
    assert a["energy"] < 1e-5 and b["energy"] < a["energy"] / 20.0


def test_the_coriolis_term_does_no_work():
    """sum_e dc dv hbar_e u_e (du_e/dt)_PV = 0 to round-off: the antisymmetric weights and the symmetric potential
    vorticity average of PotentialVortHAdvOnEdge (TendencyTerms.h:81-108)"""
    for case in (cases.sphere_case, cases.land_case, cases.hept_case):
        L, cfg, (h, u, tr), dt = case(3, 2, 1, **INVISCID)
        u = u.copy(); u[L.CellsOnEdge[:, 1] == L.NCellsAll] = 0.0      # no flow through the coast
        only_pv = capi.TendConfig.default(**{**INVISCID, "ThicknessFluxTendencyEnable": 0, "KETendencyEnable": 0, "SSHTendencyEnable": 0})
        o = Oracle(L, only_pv, 1)
        o.tend_all(h, u, tr)
        ne = L.NEdgesOwned
        hbar = o.aux["FluxLayerThickEdge"][:ne]
        w = (L.DcEdge[:ne] * L.DvEdge[:ne])[:, None] * hbar * u[:ne]
        ut = o.tend["NormalVelocityTend"][:ne]
        assert np.abs(ut).max() > 1e-5
        assert np.abs((w * ut).sum(0)).max() <= 1e-12 * np.abs(w * ut).sum(0).max(), case.__name__
        o.close()
