"""Stage coefficients of the C++ TimeStepper mirror (OmegaB200Stepper.h: setTimeStep, rational arithmetic as
src/infra/TimeMgr.cpp:382-391,747-766 evaluates RKA[s] * TimeStep) against capi.rk_coeffs, bit for bit, for whole,
fractional and awkward time steps.  capi.rk_coeffs itself is pinned to the reference's TimeMgr by the stepping tests
that run both with dt = 10 s (tests/test_oracle_vs_ref.py) and dt = 0.2 s / 0.1 s (tests/test_custom_tendency.py)."""
import os
import subprocess

import pytest

from omega_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "coeffs_test")


@pytest.fixture(scope="module")
def exe():
    if not os.path.exists(capi.LIB_PATH):
        pytest.skip("libomega_b200.so not built")
    lib_dir = os.path.join(ROOT, "omega_b200", "csrc")
    subprocess.run(["g++", "-std=c++17", "-O1", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "omega_b200", "host"),
                    os.path.join(ROOT, "tests", "cpp", "coeffs_test.cpp"), "-o", EXE, "-L" + lib_dir, "-lomega_b200",
                    "-L/usr/local/cuda/lib64", "-lcudart", "-Wl,-rpath," + lib_dir, "-Wl,-rpath,/usr/local/cuda/lib64"], check=True)
    return EXE


@pytest.mark.parametrize("dt", [60.0, 240.0, 10.0, 7.0, 1.0, 0.5, 0.2, 0.1, 30.5, 0.05, 1.25, 100.001, 3600.0])
@pytest.mark.parametrize("stepper", [capi.RK4, capi.RK2, capi.FB])
def test_mirror_coefficients_equal_rk_coeffs(exe, stepper, dt):
    out = subprocess.run([exe, str(stepper), repr(dt)], capture_output=True, text=True, check=True).stdout.split("\n")
    n, got_dt = out[0].split()
    want = capi.rk_coeffs(stepper, dt)
    assert int(n) == want.NStages
    assert float(got_dt) == want.Dt, (got_dt, want.Dt)
    for s in range(want.NStages):
        a, b = (float(x) for x in out[1 + s].split())
        assert a == want.A[s] and b == want.B[s], (stepper, dt, s, a, want.A[s], b, want.B[s])


@pytest.mark.parametrize("dt", [60.0, 240.0, 10.0, 7.0, 1.0, 0.5, 0.2, 0.1, 30.5, 0.05, 1.25, 100.001, 3600.0])
def test_rk_coeffs_equal_the_reference_time_arithmetic(dt):
    """Mult * TimeStep through the reference's own TimeInterval / TimeFrac code (oracle/_ref, TimeMgr.cpp compiled
    unmodified), for the multipliers its steppers use: 1/6, 1/3, 1/2, 1 as the doubles they are stored as"""
    import ctypes as C
    from tests import oracle_lib
    if not os.path.exists(oracle_lib.REF):
        pytest.skip("oracle/_ref not built")
    lib = C.CDLL(oracle_lib.REF)
    lib.ref_time_coeff.argtypes = [C.c_double, C.c_double, C.POINTER(C.c_double)]

    def ref(mult):
        out = C.c_double()
        assert lib.ref_time_coeff(dt, mult, C.byref(out)) == 0
        return out.value
    rk4, rk2 = capi.rk_coeffs(capi.RK4, dt), capi.rk_coeffs(capi.RK2, dt)
    assert rk4.Dt == ref(1.0)
    assert [rk4.A[s] for s in range(4)] == [ref(0.0), ref(0.5), ref(0.5), ref(1.0)]
    assert [rk4.B[s] for s in range(4)] == [ref(1.0 / 6), ref(1.0 / 3), ref(1.0 / 3), ref(1.0 / 6)]
    assert rk2.A[1] == ref(0.5) and rk2.B[1] == ref(1.0)


def test_mirror_classes_follow_the_ownership_convention():
    """create / get / getDefault / erase / clear of the reference's classes (Ownership) on the mirror:
    tests/cpp/registry_test.cpp, host code only"""
    if not os.path.exists(capi.LIB_PATH):
        pytest.skip("libomega_b200.so not built")
    lib_dir = os.path.join(ROOT, "omega_b200", "csrc")
    exe = os.path.join(ROOT, "tests", "cpp", "registry_test")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", "-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "omega_b200", "host"),
                    os.path.join(ROOT, "tests", "cpp", "registry_test.cpp"), "-o", exe, "-L" + lib_dir, "-lomega_b200",
                    "-L/usr/local/cuda/lib64", "-lcudart", "-Wl,-rpath," + lib_dir, "-Wl,-rpath,/usr/local/cuda/lib64"], check=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0 and "registry ok" in r.stdout, r.stderr
