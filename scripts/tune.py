"""Exploration harness: per-kernel times of the fused pipeline for several
settings of the OMEGA_* tuning knobs, one problem build.  Not part of the bench."""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from omega_b200 import analytic, capi, decomp, model  # noqa: E402

workload = sys.argv[1]
configs = json.loads(sys.argv[2])  # list of {env: value}
m, cfg, K, T, dt, ks = bench.build_problem(workload, 1)
L = decomp.build_all(m, 1, K)[0]
h, u, tr = analytic.tc5(L, T)
mesh = model.HorzMesh(L, reorder=True)
st = model.Stepper(mesh, cfg, T)
state = model.OceanState(mesh, T)
state.copyToDevice(h, u, tr)
co = capi.rk_coeffs(capi.RK4, dt)
names = ["fA", "fB", "fT2", "fC", "fE"]
for c in configs:
    for k in list(os.environ):
        if k.startswith("OMEGA_"):
            del os.environ[k]
    os.environ.update({k: str(v) for k, v in c.items()})
    for _ in range(2):
        st.doStep(state, capi.RK4, co, fused=True, compat=True)
    torch.cuda.synchronize()
    st.lib.omega_stepper_set_profiling(st.h, 1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    n = 3
    for _ in range(n):
        st.doStep(state, capi.RK4, co, fused=True, compat=True)
    e1.record()
    torch.cuda.synchronize()
    kms, kc = (C.c_double * 5)(), (C.c_int * 5)()
    st.lib.omega_stepper_get_profile(st.h, kms, kc, 5)
    st.lib.omega_stepper_set_profiling(st.h, 0)
    per = {nm: round(kms[i] / max(kc[i], 1), 3) for i, nm in enumerate(names)}
    print(json.dumps({"cfg": c, "ms_per_step": round(e0.elapsed_time(e1) / n, 2), "kernels": per}), flush=True)
