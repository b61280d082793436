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
           src, "-o", EXE, "-L" + lib_dir, "-lomega_b200", "-Wl,-rpath," + lib_dir]
    subprocess.run(cmd, check=True)


def test_host_mirror_compiles_against_the_c_abi():
    """CPU check: the mirror header and the test driver compile and link."""
    if not os.path.exists(capi.LIB_PATH):
        pytest.skip("libomega_b200.so not built")
    build_exe()
    assert os.path.exists(EXE)


@pytest.mark.gpu
@pytest.mark.parametrize("stepper,fused", [(capi.RK4, 1), (capi.RK4, 0), (capi.FB, 0)])
def test_host_mirror_matches_oracle(tmp_path, stepper, fused):
    if not os.path.exists(EXE):
        build_exe()
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
        f.write(f"NTracers {T}\nNSteps {nsteps}\nDt {dt}\nStepper {stepper}\nFused {fused}\n")
    with open(os.path.join(d, "cfg.bin"), "wb") as f:
        f.write(bytes(cfg))
    h.tofile(os.path.join(d, "h.bin")); u.tofile(os.path.join(d, "u.bin")); tr.tofile(os.path.join(d, "tr.bin"))
    subprocess.run([EXE, d], check=True)
    o = Oracle(L, cfg, T)
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
