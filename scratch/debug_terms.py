import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from omega_b200 import capi, model
from tests import cases
from tests.oracle_lib import Oracle
from tests.cases import rel_l2
for name, kw in (("wind", dict(WindForcingTendencyEnable=1)), ("drag", dict(BottomDragTendencyEnable=1, BottomDragCoeff=1.0e-3))):
  for nsteps in (1, 2):
    L, cfg, (h, u, tr), dt = cases.sphere_case(3, 16, 3, **kw)
    zonal = np.ascontiguousarray(0.1 * np.cos(L.LatCell)); zonal[-1] = 0.0
    merid = np.ascontiguousarray(0.05 * np.sin(2 * L.LonCell) * np.cos(L.LatCell)); merid[-1] = 0.0
    T = tr.shape[0]
    mesh = model.HorzMesh(L); st = model.Stepper(mesh, cfg, T); state = model.OceanState(mesh, T)
    o = Oracle(L, cfg, T)
    o.set_wind_stress(zonal, merid); st.set_wind_stress(zonal, merid)
    o.make_state(h, u, tr); state.copyToDevice(h, u, tr)
    co = capi.rk_coeffs(capi.RK4, dt)
    for _ in range(nsteps):
        o.do_step(capi.RK4, co, compat=1); st.doStep(state, capi.RK4, co, fused=True, compat=True)
    got, exp = state.copyToHost(0), o.current()
    d = np.abs(got[1] - exp[1])
    print(name, nsteps, os.environ.get("OMEGA_TILE_E"), [rel_l2(a, b) for a, b in zip(got, exp)], "max|du| per level", d.max(0)[[0,1,7,14,15]], flush=True)
    o.close(); st.close(); mesh.close()
