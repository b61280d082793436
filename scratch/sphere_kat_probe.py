"""Probe: do the reference's sphere golden norms (TestSetupSphere, Icos480 = 2562 cells) come out on the bundled
bisection icosahedral mesh?  (They need the fixture mesh; this measures how close the generator's mesh is.)"""
import sys
import numpy as np
sys.path.insert(0, ".")
from omega_b200 import capi, decomp, meshgen
from tests.oracle_lib import Oracle
from tests.test_oracle_kat import errors

R = 6371220.0
vX = lambda lon, lat: -R * np.sin(lon) ** 2 * np.cos(lat) ** 3
vY = lambda lon, lat: -4 * R * np.sin(lon) * np.cos(lon) * np.cos(lat) ** 3 * np.sin(lat)
div = lambda lon, lat: np.sin(lon) * np.cos(lon) * np.cos(lat) ** 2 * (20 * np.sin(lat) ** 2 - 6)
scal = lambda lon, lat: R * np.cos(lon) * np.cos(lat) ** 4
gX = lambda lon, lat: -np.sin(lon) * np.cos(lat) ** 3
gY = lambda lon, lat: -4 * np.cos(lon) * np.cos(lat) ** 3 * np.sin(lat)
EXP = {"Div": (0.0136595773989796766, 0.00367052484586384131), "Grad": (0.00187912292540623471, 0.00149841802817334935)}

level = int(sys.argv[1]) if len(sys.argv) > 1 else 4
m = meshgen.icosahedral(level, radius=R)
K = 2
L = decomp.build_all(m, 1, K)[0]
off = dict(ThicknessFluxTendencyEnable=0, PVTendencyEnable=0, KETendencyEnable=0, SSHTendencyEnable=0, VelDiffTendencyEnable=0,
           VelHyperDiffTendencyEnable=0, TracerHorzAdvTendencyEnable=0, TracerDiffTendencyEnable=0, TracerHyperDiffTendencyEnable=0)
o = Oracle(L, capi.TendConfig.default(**off), 0)
rep = lambda a: np.ascontiguousarray(np.repeat(a[:, None], K, 1))
un = np.cos(L.AngleEdge) * vX(L.LonEdge, L.LatEdge) + np.sin(L.AngleEdge) * vY(L.LonEdge, L.LatEdge)
u = rep(un); u[-1] = 0
o.set_config(capi.TendConfig.default(**{**off, "ThicknessFluxTendencyEnable": 1}))
o.aux["FluxLayerThickEdge"][:] = 1.0
o.tend_thickness_only(u)
ex = rep(-div(L.LonCell, L.LatCell)); ex[-1] = 0
print("Div ", errors(o.tend["LayerThicknessTend"][:-1], ex[:-1], L.AreaCell[:-1]), "expected", EXP["Div"])
o.set_config(capi.TendConfig.default(**{**off, "KETendencyEnable": 1}))
ke = rep(scal(L.LonCell, L.LatCell)); ke[-1] = 0
o.aux["KineticEnergyCell"][:] = ke
o.tend_velocity_only(u)
gn = rep(-(np.cos(L.AngleEdge) * gX(L.LonEdge, L.LatEdge) + np.sin(L.AngleEdge) * gY(L.LonEdge, L.LatEdge))); gn[-1] = 0
aE = (L.DcEdge * L.DvEdge / 2)[:-1]
print("Grad", errors(o.tend["NormalVelocityTend"][:-1], gn[:-1], aE), "expected", EXP["Grad"])
