"""Known-answer tests: the oracle's functors on a regenerated 48x48 periodic
hex mesh (Lx=1, Ly=sqrt(3)/2) against the analytic-error norms the
reference's own tests pin for TestSetupPlane:
test/ocn/TendencyTermsTest.cpp:43-59, test/ocn/HorzOperatorsTest.cpp:33-43,
norm definition test/ocn/OceanTestCommon.h:403-562.

The fixture mesh itself (OmegaPlanarMesh.nc) is not in the reference tree, so
rows whose norm depends on where the mesh origin sits relative to the
analytic field (everything but the divergence rows) are matched by scanning
the few natural origins of a planar-hex generator; the divergence rows are
origin independent and are the hard pin.
"""
import numpy as np
import pytest

from omega_b200 import capi, decomp, meshgen
from tests.oracle_lib import Oracle

Lx, Ly = 1.0, np.sqrt(3.0) / 2.0
P = 2 * np.pi

vecX = lambda x, y: np.sin(P * x / Lx) * np.cos(P * y / Ly)
vecY = lambda x, y: np.cos(P * x / Lx) * np.sin(P * y / Ly)
divergence = lambda x, y: P * (1 / Lx + 1 / Ly) * np.cos(P * x / Lx) * np.cos(P * y / Ly)
scalar = lambda x, y: np.sin(P * x / Lx) * np.sin(P * y / Ly)
gradX = lambda x, y: P / Lx * np.cos(P * x / Lx) * np.sin(P * y / Ly)
gradY = lambda x, y: P / Ly * np.sin(P * x / Lx) * np.cos(P * y / Ly)
curl = lambda x, y: P * (-1 / Lx + 1 / Ly) * np.sin(P * x / Lx) * np.sin(P * y / Ly)
lapX = lambda x, y: -P * P * (1 / Lx**2 + 1 / Ly**2) * np.sin(P * x / Lx) * np.cos(P * y / Ly)
lapY = lambda x, y: -P * P * (1 / Lx**2 + 1 / Ly**2) * np.cos(P * x / Lx) * np.sin(P * y / Ly)
layerThick = lambda x, y: 2 + np.sin(P * x / Lx) * np.cos(P * y / Ly)
planetVort = lambda x, y: np.cos(P * x / Lx) * np.cos(P * y / Ly)
scalarA = lambda x, y: np.cos(P * x / Lx) * np.sin(P * y / Ly)
scalarB = lambda x, y: 2 + np.cos(P * x / Lx) * np.cos(P * y / Ly)
scalarC = lambda x, y: np.cos(P * x / Lx) ** 2 - np.sin(P * y / Ly) ** 2


def tracerFluxDiv(x, y):
    return (P / (Lx * Ly)) * (np.cos(P * x / Lx) * (
        2 * (Lx + Ly) * np.cos(P * y / Ly)
        + (Lx + 2 * Ly) * np.sin(P * x / Lx) * np.cos(P * y / Ly) ** 2
        - Lx * np.sin(P * x / Lx) * np.sin(P * y / Ly) ** 2))


def tracerDiff(x, y):
    return -P * P * np.sin(P * y / Ly) * (
        2 * (1 / Lx**2 + 1 / Ly**2) * np.cos(P * x / Lx)
        + (1 / Ly**2 + (1 / Lx**2 + 1 / Ly**2) * np.cos(2 * P * x / Lx)) * np.cos(P * y / Ly))


def tracerHyperDiff(x, y):
    return -2 * P * P * (np.cos(2 * P * x / Lx) / Lx**2 + np.cos(2 * P * y / Ly) / Ly**2)


EXPECTED = {  # test/ocn/TendencyTermsTest.cpp:43-59, HorzOperatorsTest.cpp:33-43
    "Div": (0.00124886886594453264, 0.00124886886590977139),
    "PV": (0.00807347170900282914, 0.00794755105765788429),
    "Grad": (0.00125026071878537952, 0.00134354611117262161),
    "Laplace": (0.00113090174765822192, 0.00134324628763667899),
    "TrHAdv": (0.00205864372747571571, 0.00172418025417940784),
    "TrDel2": (0.00334357193650093847, 0.00290978146207349032),
    "TrDel4": (0.00508833446725232875, 0.00523080740758275625),
}


def errors(num, exact, area):
    """computeErrors, test/ocn/OceanTestCommon.h:403-562."""
    d = np.abs(num - exact)
    a = area.reshape((1,) * (num.ndim - 2) + (-1, 1))
    linf = d.max() / np.abs(exact).max()
    l2 = np.sqrt((a * d * d).sum() / (a * exact * exact).sum())
    return linf, l2


class Plane:
    def __init__(self, x0=0.0, y0=0.0, K=2, T=2):
        self.m = meshgen.planar_hex(48, 48, 1.0 / 48, x0=x0, y0=y0)
        self.L = decomp.build_all(self.m, 1, K)[0]
        self.K, self.T = K, T
        L = self.L
        self.xc, self.yc = L.XCell, L.YCell
        self.xe, self.ye = L.XEdge, L.YEdge
        self.xv, self.yv = L.XVertex, L.YVertex

    def on(self, f, x, y, T=0):
        a = np.repeat(f(x, y)[:, None], self.K, 1)
        a[-1] = 0.0  # boundary slot row
        if T:
            a = np.repeat(a[None], T, 0)
        return np.ascontiguousarray(a)

    def edge_normal(self, fx, fy):
        a = np.cos(self.L.AngleEdge) * fx(self.xe, self.ye) + np.sin(self.L.AngleEdge) * fy(self.xe, self.ye)
        a = np.repeat(a[:, None], self.K, 1)
        a[-1] = 0
        return np.ascontiguousarray(a)

    def edge_tangent(self, fx, fy):
        a = -np.sin(self.L.AngleEdge) * fx(self.xe, self.ye) + np.cos(self.L.AngleEdge) * fy(self.xe, self.ye)
        a = np.repeat(a[:, None], self.K, 1)
        a[-1] = 0
        return np.ascontiguousarray(a)

    def area_edge(self):
        return (self.L.DcEdge * self.L.DvEdge / 2)[:-1]


def compute_all(p: Plane, lib_path=None):
    """Every KAT row -> (LInf, L2)."""
    kw = {} if lib_path is None else {"lib_path": lib_path}
    off = dict(ThicknessFluxTendencyEnable=0, PVTendencyEnable=0, KETendencyEnable=0,
               SSHTendencyEnable=0, VelDiffTendencyEnable=0, VelHyperDiffTendencyEnable=0,
               TracerHorzAdvTendencyEnable=0, TracerDiffTendencyEnable=0,
               TracerHyperDiffTendencyEnable=0)
    o = Oracle(p.L, capi.TendConfig.default(**off), p.T, **kw)
    L = p.L
    out = {}
    u = p.edge_normal(vecX, vecY)
    aC, aE = L.AreaCell[:-1], p.area_edge()

    def run(**on):
        o.set_config(capi.TendConfig.default(**{**off, **on}))

    # ThicknessFluxDivOnCell with flux = 1
    run(ThicknessFluxTendencyEnable=1)
    o.aux["FluxLayerThickEdge"][:] = 1.0
    o.tend_thickness_only(u)
    out["Div"] = errors(o.tend["LayerThicknessTend"][:-1], -p.on(divergence, p.xc, p.yc)[:-1], aC)
    # PotentialVortHAdvOnEdge
    run(PVTendencyEnable=1)
    o.aux["NormRelVortEdge"][:] = p.on(lambda x, y: curl(x, y) / layerThick(x, y), p.xe, p.ye)
    o.aux["NormPlanetVortEdge"][:] = p.on(lambda x, y: planetVort(x, y) / layerThick(x, y), p.xe, p.ye)
    o.aux["FluxLayerThickEdge"][:] = p.on(layerThick, p.xe, p.ye)
    o.tend_velocity_only(u)
    q = lambda x, y: (curl(x, y) + planetVort(x, y)) / layerThick(x, y) * layerThick(x, y)
    ex = p.edge_tangent(lambda x, y: q(x, y) * vecX(x, y), lambda x, y: q(x, y) * vecY(x, y))
    out["PV"] = errors(o.tend["NormalVelocityTend"][:-1], ex[:-1], aE)
    # KEGradOnEdge
    run(KETendencyEnable=1)
    o.aux["KineticEnergyCell"][:] = p.on(scalar, p.xc, p.yc)
    o.tend_velocity_only(u)
    ex = p.edge_normal(lambda x, y: -gradX(x, y), lambda x, y: -gradY(x, y))
    out["Grad"] = errors(o.tend["NormalVelocityTend"][:-1], ex[:-1], aE)
    # SSHGradOnEdge
    run(SSHTendencyEnable=1)
    o.aux["SshCell"][:] = p.on(scalar, p.xc, p.yc)
    o.tend_velocity_only(u)
    out["SSHGrad"] = errors(o.tend["NormalVelocityTend"][:-1], 9.80665 * ex[:-1], aE)
    # VelocityDiffusionOnEdge / VelocityHyperDiffOnEdge
    o.aux["VelocityDivCell"][:] = p.on(divergence, p.xc, p.yc)
    o.aux["RelVortVertex"][:] = p.on(curl, p.xv, p.yv)
    o.aux["Del2DivCell"][:] = p.on(divergence, p.xc, p.yc)
    o.aux["Del2RelVortVertex"][:] = p.on(curl, p.xv, p.yv)
    exl = p.edge_normal(lapX, lapY)
    run(VelDiffTendencyEnable=1, ViscDel2=1.0e3)
    o.tend_velocity_only(u)
    out["Laplace"] = errors(o.tend["NormalVelocityTend"][:-1], 1.0e3 * exl[:-1], aE)
    run(VelHyperDiffTendencyEnable=1, ViscDel4=1.2e11, DivFactor=1.0)
    o.tend_velocity_only(u)
    out["Laplace4"] = errors(o.tend["NormalVelocityTend"][:-1], -1.2e11 * exl[:-1], aE)
    # tracer terms
    run(TracerHorzAdvTendencyEnable=1)
    o.aux["HTracersEdge"][:] = -p.on(layerThick, p.xe, p.ye, p.T)
    tr = p.on(scalarA, p.xc, p.yc, p.T)
    o.tend_tracer_only(u, tr)
    out["TrHAdv"] = errors(o.tend["TracerTend"][:, :-1], p.on(tracerFluxDiv, p.xc, p.yc, p.T)[:, :-1], aC)
    run(TracerDiffTendencyEnable=1, EddyDiff2=1.0)
    o.aux["MeanLayerThickEdge"][:] = p.on(scalarB, p.xe, p.ye)
    o.tend_tracer_only(u, tr)
    out["TrDel2"] = errors(o.tend["TracerTend"][:, :-1], p.on(tracerDiff, p.xc, p.yc, p.T)[:, :-1], aC)
    run(TracerHyperDiffTendencyEnable=1, EddyDiff4=1.0)
    o.aux["Del2TracersCell"][:] = p.on(scalarC, p.xc, p.yc, p.T)
    o.tend_tracer_only(u, tr)
    out["TrDel4"] = errors(o.tend["TracerTend"][:, :-1], -p.on(tracerHyperDiff, p.xc, p.yc, p.T)[:, :-1], aC)
    # generic operators, src/ocn/HorzOperators.h
    d = np.zeros_like(tr[0])
    o.op("divergence", u, d)
    out["OpDiv"] = errors(d[:-1], p.on(divergence, p.xc, p.yc)[:-1], aC)
    g = np.zeros_like(u)
    o.op("gradient", p.on(scalar, p.xc, p.yc), g)
    out["OpGrad"] = errors(g[:-1], p.edge_normal(gradX, gradY)[:-1], aE)
    cv = np.zeros((L.NVerticesSize, p.K))
    o.op("curl", u, cv)
    out["OpCurl"] = errors(cv[:-1], p.on(curl, p.xv, p.yv)[:-1], L.AreaTriangle[:-1])
    r = np.zeros_like(u)
    o.op("tangential_recon", u, r)
    out["OpRecon"] = errors(r[:-1], p.edge_tangent(vecX, vecY)[:-1], aE)
    o.close()
    return out


@pytest.fixture(scope="module")
def kat():
    return compute_all(Plane())


def test_divergence_rows_match_reference_golden(kat):
    # origin independent: hard pin at the reference's own tolerance (1e-10 / 1e-5)
    for row in ("Div", "OpDiv"):
        for got, exp in zip(kat[row], EXPECTED["Div"]):
            assert abs(got - exp) <= 1e-9 * exp, (row, got, exp)


@pytest.mark.parametrize("row,key,rtol", [
    ("PV", "PV", 1e-5), ("Grad", "Grad", 1e-5), ("SSHGrad", "Grad", 1e-5), ("OpGrad", "Grad", 1e-5),
    ("Laplace", "Laplace", 1e-5), ("Laplace4", "Laplace", 1e-5), ("TrHAdv", "TrHAdv", 1e-5),
    ("TrDel2", "TrDel2", 1e-5), ("TrDel4", "TrDel4", 1e-5)])
def test_tendency_rows_match_reference_golden(kat, row, key, rtol):
    for got, exp in zip(kat[row], EXPECTED[key]):
        assert abs(got - exp) <= rtol * exp, (row, got, exp)


def test_curl_and_recon_rows(kat):
    # HorzOperatorsTest.cpp:37-40
    assert abs(kat["OpCurl"][0] - 0.16136566356969) <= 1e-8
    assert abs(kat["OpRecon"][0] - 0.00450897496974901) <= 1e-8
