"""Small fused RK4 runs for compute-sanitizer (memcheck / racecheck): K = 16 so the tiled kernels run,
T = 1, 2, 3, 5 tracers (short and long copy schedules of k_fCT2; OMEGA_TILE_V2=0 selects the older
k_fCT with its named-barrier protocol for T >= 2), two steps each, plus one 2-rank emulated exchange
round through the round pack / unpack kernels.  usage: compute-sanitizer --tool X python scripts/sanitize_case.py"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from omega_b200 import capi, model  # noqa: E402
from tests import cases  # noqa: E402

v2 = os.environ.get("OMEGA_TILE_V2", "1") != "0"
for T in (1, 2, 3, 5):
    if not v2 and T < 2:
        continue
    L, cfg, (h, u, tr), dt = cases.sphere_case(3, 16, T)
    mesh = model.HorzMesh(L, reorder=True)
    st, state = model.Stepper(mesh, cfg, T), model.OceanState(mesh, T)
    state.copyToDevice(h, u, tr)
    co = capi.rk_coeffs(capi.RK4, dt)
    for _ in range(2):
        st.doStep(state, capi.RK4, co, fused=True, compat=True)
    torch.cuda.synchronize()
    gh, gu, gt = state.copyToHost(0)
    assert np.isfinite(gh).all() and np.isfinite(gu).all() and np.isfinite(gt).all()
    print("T", T, "path bits", st.last_path(), "launches", st.launch_count(), flush=True)
    st.close(); mesh.close()
print("SANITIZE_CASE_DONE", flush=True)
