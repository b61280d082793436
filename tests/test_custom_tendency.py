"""Host custom tendencies (Tendencies::CustomTendencyType, src/ocn/Tendencies.h:51-53) and the reference's own
TimeStepperTest (test/timeStepping/TimeStepperTest.cpp): every built-in term off, du/dt = -0.5 u through a custom
velocity tendency, h = u = tracer = 1, steps of 0.2 s and 0.1 s to t = 1 s, convergence order 4 (RK4), 1
(forward-backward), 2 (RK2) within 0.1 -- on the reference's own stepper sources (oracle/_ref), on the C oracle (equal
to them bit for bit), and on the GPU through omega_stepper_set_custom_tendencies (the step then runs on the
kernel-per-term path with the callback in every stage).  The fractional steps go through the reference's rational
time arithmetic."""
import os

import numpy as np
import pytest

from omega_b200 import capi
from tests import cases, oracle_lib
from tests.oracle_lib import Oracle

COEFF, T_END = 0.5, 1.0
ALL_OFF = dict(ThicknessFluxTendencyEnable=0, PVTendencyEnable=0, KETendencyEnable=0, SSHTendencyEnable=0, VelDiffTendencyEnable=0,
               VelHyperDiffTendencyEnable=0, TracerHorzAdvTendencyEnable=0, TracerDiffTendencyEnable=0,
               TracerHyperDiffTendencyEnable=0, WindForcingTendencyEnable=0, BottomDragTendencyEnable=0)
EXPECTED_ORDER = {capi.RK4: 4.0, capi.FB: 1.0, capi.RK2: 2.0}      # TimeStepperTest.cpp:375-388


def decay(tend, h, u, time):
    """DecayVelocityTendency, TimeStepperTest.cpp:49-72"""
    tend -= COEFF * u


def problem(K=1, T=1):
    L, cfg, _, _ = cases.sphere_case(2, K, T, **ALL_OFF)
    h = np.ones((L.NCellsSize, K)); u = np.ones((L.NEdgesSize, K)); tr = np.ones((T, L.NCellsSize, K))
    h[-1] = 0; u[-1] = 0; tr[:, -1] = 0
    return L, cfg, h, u, tr


def run_oracle(lib_path, stepper, dt):
    L, cfg, h, u, tr = problem()
    o = Oracle(L, cfg, 1, lib_path=lib_path)
    o.set_custom_tendencies(None, decay)
    o.make_state(h, u, tr)
    n = int(round(T_END / dt))
    for s in range(n):
        o.do_step(stepper, capi.rk_coeffs(stepper, dt, t0=s * dt))
    out = [a.copy() for a in o.current()]
    o.close()
    return L, out


def linf_error(L, state):
    """computeErrors of the test (TimeStepperTest.cpp:140-190): velocity against exp(-Coeff t) on the edges"""
    return float(np.abs(state[1][:L.NEdgesAll] - np.exp(-COEFF * T_END)).max())


@pytest.mark.parametrize("stepper", [capi.RK4, capi.FB, capi.RK2])
def test_reference_timestepper_test_on_the_reference_and_the_oracle(stepper):
    if not os.path.exists(oracle_lib.REF):
        pytest.skip("oracle/_ref not built")
    errs = []
    for dt in (0.2, 0.1):
        L, ref = run_oracle(oracle_lib.REF, stepper, dt)
        _, mine = run_oracle(oracle_lib.STRICT, stepper, dt)
        for a, b in zip(ref, mine):
            assert np.array_equal(a, b), (stepper, dt)
        assert np.array_equal(ref[0][:-1], np.ones_like(ref[0][:-1])) and np.array_equal(ref[2][:, :-1], np.ones_like(ref[2][:, :-1]))
        errs.append(linf_error(L, ref))
    rate = np.log2(errs[0] / errs[1])
    assert abs(rate - EXPECTED_ORDER[stepper]) <= 0.1, (stepper, errs, rate)


def test_custom_terms_are_added_after_the_built_in_ones():
    """one right-hand side with every Default.yml term on: the custom thickness and velocity terms land on top of the
    built-in tendencies (Tendencies.cpp:288-293, 416-421), in the oracle and in the reference build alike"""
    L, cfg, (h, u, tr), dt = cases.sphere_case(3, 4, 2)
    seen = []

    def thick(tend, hh, uu, time):
        seen.append(("h", hh is not None, time))
        tend += 1.0e-3

    def vel(tend, hh, uu, time):
        seen.append(("u", hh is not None, time))
        tend -= COEFF * uu

    libs = [oracle_lib.STRICT] + ([oracle_lib.REF] if os.path.exists(oracle_lib.REF) else [])
    res = []
    for lp in libs:
        o = Oracle(L, cfg, 2, lib_path=lp)
        o.tend_all(h, u, tr, 30.0)
        base = {k: v.copy() for k, v in o.tend.items()}
        o.set_custom_tendencies(thick, vel)
        o.tend_all(h, u, tr, 30.0)
        nC, nE = L.NCellsAll, L.NEdgesAll
        assert np.array_equal(o.tend["LayerThicknessTend"][:nC], base["LayerThicknessTend"][:nC] + 1.0e-3)
        assert np.array_equal(o.tend["NormalVelocityTend"][:nE], base["NormalVelocityTend"][:nE] - COEFF * u[:nE])
        assert np.array_equal(o.tend["TracerTend"], base["TracerTend"])
        res.append({k: v.copy() for k, v in o.tend.items()})
        o.set_custom_tendencies(None, None)
        o.tend_all(h, u, tr, 30.0)
        assert all(np.array_equal(o.tend[k], base[k]) for k in base)
        o.close()
    assert ("h", True, 30.0) in seen and ("u", True, 30.0) in seen
    if len(res) == 2:
        assert all(np.array_equal(res[0][k], res[1][k]) for k in res[0])


@pytest.mark.gpu
@pytest.mark.parametrize("stepper", [capi.RK4, capi.FB, capi.RK2])
def test_reference_timestepper_test_on_the_gpu(stepper):
    """the same test through the library: torch ops on the device arrays as the custom term, Fused=1 requested"""
    from omega_b200 import model
    errs = []
    for dt in (0.2, 0.1):
        L, cfg, h, u, tr = problem()
        mesh = model.HorzMesh(L, reorder=True)
        st, state = model.Stepper(mesh, cfg, 1), model.OceanState(mesh, 1)
        st.set_custom_tendencies(None, decay)
        state.copyToDevice(h, u, tr)
        n = int(round(T_END / dt))
        for s in range(n):
            st.doStep(state, stepper, capi.rk_coeffs(stepper, dt, t0=s * dt), fused=True, compat=True)
        assert st._custom_error is None, st._custom_error
        assert st.last_path() & 1 == 0, "a custom term must take the step off the fused kernels"
        got = state.copyToHost(0)
        _, want = run_oracle(oracle_lib.STRICT, stepper, dt)
        for a, b in zip(got, want):
            assert cases.rel_l2(a, b) <= 1e-12
        errs.append(linf_error(L, got))
        st.close(); mesh.close()
    rate = np.log2(errs[0] / errs[1])
    assert abs(rate - EXPECTED_ORDER[stepper]) <= 0.1, (stepper, errs, rate)


@pytest.mark.gpu
def test_custom_terms_on_top_of_every_built_in_term_on_the_gpu():
    from omega_b200 import model
    L, cfg, (h, u, tr), dt = cases.sphere_case(3, 16, 2)

    # rates that the quarter-hour steps of this case resolve (the decay of the TimeStepperTest is 0.5 / s)
    def thick(tend, hh, uu, time):
        tend += 1.0e-7 * hh

    def vel(tend, hh, uu, time):
        tend -= 1.0e-5 * uu

    o = Oracle(L, cfg, 2)
    o.set_custom_tendencies(thick, vel)
    o.make_state(h, u, tr)
    mesh = model.HorzMesh(L, reorder=True)
    st, state = model.Stepper(mesh, cfg, 2), model.OceanState(mesh, 2)
    st.set_custom_tendencies(thick, vel)
    state.copyToDevice(h, u, tr)
    for s in range(3):
        co = capi.rk_coeffs(capi.RK4, dt, t0=s * dt)
        o.do_step(capi.RK4, co)
        st.doStep(state, capi.RK4, co, fused=True, compat=True)
    assert st._custom_error is None, st._custom_error
    plain = Oracle(L, cfg, 2)
    plain.make_state(h, u, tr)
    for s in range(3):
        plain.do_step(capi.RK4, capi.rk_coeffs(capi.RK4, dt, t0=s * dt))
    for a, b, c in zip(state.copyToHost(0), o.current(), plain.current()):
        assert np.isfinite(b).all() and cases.rel_l2(a, b) <= 1e-10
    assert cases.rel_l2(plain.current()[1], o.current()[1]) > 1e-4, "the custom terms must matter"
    plain.close()
    # removed again: the fused kernels are back
    st.set_custom_tendencies(None, None)
    st.doStep(state, capi.RK4, capi.rk_coeffs(capi.RK4, dt, t0=3 * dt), fused=True, compat=True)
    assert st.last_path() & 1 == 1
    o.close(); st.close(); mesh.close()
