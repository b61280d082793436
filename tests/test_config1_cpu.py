"""BASELINE configs[0] on the CPU oracle: planar hex, one level, manufactured solution, RK4.

Pins the oracle's (and the compiled reference's) manufactured source terms
(src/ocn/CustomTendencyTerms.cpp:107-210) and the mesh generator against the
exact solution: the error after a quarter period converges at second order in
space (doc/design/OmegaV0ShallowWater.md:656-666).  The GPU twin of this test is
tests/test_gpu_parity.py::test_config1_manufactured_rk4_stepping_and_convergence."""
import os

import numpy as np
import pytest

from omega_b200 import analytic, capi, decomp, meshgen
from tests import oracle_lib
from tests.oracle_lib import Oracle


def config1_problem(nx):
    dc = 5.0e6 / nx
    m = meshgen.planar_hex(nx, nx, dc, f0=1.0e-4, bottom_depth=1000.0)
    L = decomp.build_all(m, 1, 1)[0]
    cfg = capi.TendConfig.default(ManufacturedSolutionTendency=1, ManufacturedWavelengthX=float(L.Lx),
                                  ManufacturedWavelengthY=float(L.Ly), TracerHorzAdvTendencyEnable=0,
                                  TracerDiffTendencyEnable=0, TracerHyperDiffTendencyEnable=0)
    h, u = analytic.manufactured(L, cfg)
    return L, cfg, h, u


def run(nx, dt, t_end, lib_path=oracle_lib.STRICT):
    L, cfg, h, u = config1_problem(nx)
    o = Oracle(L, cfg, 0, lib_path=lib_path)
    o.make_state(h, u, np.zeros((0, h.shape[0], 1)))
    n = int(round(t_end / dt))
    for i in range(n):
        o.do_step(capi.RK4, capi.rk_coeffs(capi.RK4, dt, t0=i * dt), compat=1)
    eh, eu, _ = (np.array(a) for a in o.current())
    o.close()
    xh, xu = analytic.manufactured(L, cfg, t=n * dt)
    a = cfg.ManufacturedAmplitude
    return L, eh, eu, (np.sqrt(np.mean((eh[:-1] - xh[:-1]) ** 2)) / a, np.sqrt(np.mean((eu[:-1] - xu[:-1]) ** 2)) / a)


def test_manufactured_solution_second_order_convergence():
    errs = [run(nx, dt, 9600.0)[3] for nx, dt in ((16, 480.0), (32, 240.0), (64, 120.0))]
    rates = [np.log2(np.array(errs[i]) / np.array(errs[i + 1])) for i in range(2)]
    assert errs[-1][0] < 5e-3 and errs[-1][1] < 5e-3, errs
    assert all(abs(r - 2.0) <= 0.35 for r in rates[-1]), (errs, rates)


def test_manufactured_solution_volume_conserved():
    L, eh, _, _ = run(32, 240.0, 4800.0)
    H0, A = float(L.BottomDepth[0]), L.AreaCell[:-1]
    assert abs((A * (eh[:-1, 0] - H0)).sum()) <= 1e-9 * (A * H0).sum()


@pytest.mark.skipif(not os.path.exists(oracle_lib.REF), reason="oracle/_ref not built (needs /root/reference)")
def test_manufactured_stepping_oracle_equals_compiled_reference():
    """the C restatement against the reference's own sources, config-1 shape, 20 RK4 steps: bit-equal"""
    _, h1, u1, _ = run(16, 480.0, 9600.0)
    _, h2, u2, _ = run(16, 480.0, 9600.0, lib_path=oracle_lib.REF)
    assert np.array_equal(h1, h2) and np.array_equal(u1, u2)
