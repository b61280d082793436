"""Pins the plain-C oracle (oracle/omega_oracle.c) bit-for-bit against
oracle/_ref/libomega_ref.so = the reference's OWN hot-path sources
(TendencyTerms.*, auxiliaryVars/*, HorzOperators.*, AuxiliaryState.cpp,
Tendencies.cpp, CustomTendencyTerms.cpp, OceanState.cpp, TimeStepper.cpp,
RungeKutta4/2Stepper.cpp, ForwardBackwardStepper.cpp, TimeMgr.cpp) compiled
unmodified over a host Kokkos stand-in (oracle/ref_shim/).  Both builds use
-ffp-contract=off, so equality is exact, including the raw time-level indexing
the reference applies on odd steps and the rational-time
stage coefficients (P2)."""
import os

import numpy as np
import pytest

from omega_b200 import analytic, capi
from tests import cases, oracle_lib
from tests.oracle_lib import Oracle

pytestmark = pytest.mark.skipif(not os.path.exists(oracle_lib.REF),
                                reason="oracle/_ref not built (needs /root/reference; see oracle/ref_shim/Makefile)")


def pair(L, cfg, T):
    return Oracle(L, cfg, T), Oracle(L, cfg, T, lib_path=oracle_lib.REF)


CASES = {
    "sphere3_K4_T2": lambda: cases.sphere_case(3, 4, 2),
    "sphere2_K3_T1": lambda: cases.sphere_case(2, 3, 1),
    "plane12_K5_T2": lambda: cases.plane_case(12, 5, 2),
    "plane16_K1_T0": lambda: cases.plane_case(16, 1, 0),
    "land3_K4_T2": lambda: cases.land_case(3, 4, 2),      # coastline: one-cell edges, EdgeMask = 0, boundary-slot reads
    "hept3_K4_T2": lambda: cases.hept_case(3, 4, 2),      # maxEdges = 7, rows of 11 and 12 edgesOnEdge
}


@pytest.mark.parametrize("name", list(CASES))
def test_rhs_bit_exact(name):
    L, cfg, (h, u, tr), dt = CASES[name]()
    a, b = pair(L, cfg, tr.shape[0])
    a.tend_all(h, u, tr, 0.0)
    b.tend_all(h, u, tr, 0.0)
    for n in capi.AUX_FIELDS:
        assert np.array_equal(a.aux[n], b.aux[n]), n
    for n in a.tend:
        assert np.array_equal(a.tend[n], b.tend[n]), n
    for x, y in zip(a.derived(), b.derived()):
        assert np.array_equal(x, y)
    a.close(); b.close()


@pytest.mark.parametrize("land", [False, True])
def test_optional_terms_bit_exact(land):
    L, cfg, (h, u, tr), dt, (zonal, merid) = cases.all_terms_case(land=land)
    a, b = pair(L, cfg, tr.shape[0])
    for o in (a, b):
        o.set_wind_stress(zonal, merid)
        o.tend_all(h, u, tr)
    for n in capi.AUX_FIELDS:
        assert np.array_equal(a.aux[n], b.aux[n]), n
    for n in a.tend:
        assert np.array_equal(a.tend[n], b.tend[n]), n
    assert np.abs(b.aux["NormalStressEdge"]).max() > 0
    a.close(); b.close()


def test_manufactured_terms_bit_exact():
    L, cfg, _, dt = cases.plane_case(16, 1, 0, ManufacturedSolutionTendency=1,
                                     ManufacturedWavelengthX=16 * 5.0e6 / 64,
                                     ManufacturedWavelengthY=16 * 5.0e6 / 64 * np.sqrt(3) / 2)
    h, u = analytic.manufactured(L, cfg)
    tr = np.zeros((0, h.shape[0], 1))
    a, b = pair(L, cfg, 0)
    a.tend_all(h, u, tr, 777.0)
    b.tend_all(h, u, tr, 777.0)
    for n in ("LayerThicknessTend", "NormalVelocityTend"):
        assert np.array_equal(a.tend[n], b.tend[n]), n
    a.close(); b.close()


@pytest.mark.parametrize("stepper", [capi.RK4, capi.RK2, capi.FB])
@pytest.mark.parametrize("name,dt_override", [("sphere3_K4_T2", None), ("plane12_K5_T2", 10.0), ("land3_K4_T2", None), ("hept3_K4_T2", None)])
def test_steppers_bit_exact(name, stepper, dt_override):
    """5 steps: odd steps exercise the raw time-level indexing; dt = 10 s makes
    RKB*dt = 10/6, 10/3 go through the reference's rational time arithmetic."""
    L, cfg, (h, u, tr), dt = CASES[name]()
    dt = dt_override or dt
    a, b = pair(L, cfg, tr.shape[0])
    for o in (a, b):
        o.make_state(h, u, tr)
    for s in range(5):
        co = capi.rk_coeffs(stepper, dt, t0=s * dt)
        a.do_step(stepper, co, compat=1)
        b.do_step(stepper, co, compat=1)
        for x, y in zip(a.current(), b.current()):
            assert np.array_equal(x, y), (name, stepper, s)
    assert np.isfinite(a.current()[0]).all()
    a.close(); b.close()


def test_manufactured_stepping_bit_exact():
    L, cfg, _, dt = cases.plane_case(16, 1, 0, ManufacturedSolutionTendency=1,
                                     ManufacturedWavelengthX=16 * 5.0e6 / 64,
                                     ManufacturedWavelengthY=16 * 5.0e6 / 64 * np.sqrt(3) / 2)
    h, u = analytic.manufactured(L, cfg)
    tr = np.zeros((0, h.shape[0], 1))
    a, b = pair(L, cfg, 0)
    for o in (a, b):
        o.make_state(h, u, tr)
    for s in range(3):
        co = capi.rk_coeffs(capi.RK4, dt, t0=s * dt)
        a.do_step(capi.RK4, co, compat=1)
        b.do_step(capi.RK4, co, compat=1)
    for x, y in zip(a.current()[:2], b.current()[:2]):
        assert np.array_equal(x, y)
    a.close(); b.close()


def test_reference_build_reproduces_its_own_golden_norms():
    """the reference's functors, run through the reference build, on the
    regenerated 48x48 plane mesh (same rows as test_oracle_kat.py)."""
    from tests import test_oracle_kat as kat
    got = kat.compute_all(kat.Plane(), lib_path=oracle_lib.REF)
    for row in ("Div", "OpDiv"):
        for g, e in zip(got[row], kat.EXPECTED["Div"]):
            assert abs(g - e) <= 1e-9 * e
    for row, key in (("PV", "PV"), ("Grad", "Grad"), ("Laplace", "Laplace"), ("TrHAdv", "TrHAdv"),
                     ("TrDel2", "TrDel2"), ("TrDel4", "TrDel4")):
        for g, e in zip(got[row], kat.EXPECTED[key]):
            assert abs(g - e) <= 1e-5 * e, row
