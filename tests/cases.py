"""Shared small test cases: mesh + state + config (host order)."""
from __future__ import annotations

import functools

import numpy as np

from omega_b200 import analytic, capi, decomp, meshgen


def rel_max(a, b):
    """BASELINE.json parity metric: max|a-b| / max|b|."""
    if b.size == 0:
        return 0.0
    s = np.abs(b).max()
    d = np.abs(a - b).max()
    return d / s if s > 0 else d


def rel_l2(a, b):
    s = np.sqrt((b * b).sum())
    return np.sqrt(((a - b) ** 2).sum()) / s if s > 0 else np.sqrt(((a - b) ** 2).sum())


@functools.lru_cache(maxsize=None)
def sphere_mesh(level):
    m = meshgen.icosahedral(level)
    m.bottomDepth = analytic.tc5_bottom_depth(m)
    return m


@functools.lru_cache(maxsize=None)
def plane_mesh(nx, ny, dc):
    return meshgen.planar_hex(nx, ny, dc, f0=1.0e-4, bottom_depth=1000.0)


def sphere_case(level, K, T, **cfgkw):
    """TC5-style state on an icosahedral mesh, every Default.yml term on plus
    tracer hyperdiffusion, coefficients scaled to the coarse test meshes."""
    m = sphere_mesh(level)
    L = decomp.build_all(m, 1, K)[0]
    dc = float(np.mean(m.dcEdge))
    kw = dict(ViscDel2=1.0e-4 * dc**2 / 60.0 * 10, ViscDel4=1.0e-3 * dc**4 / 600.0,
              EddyDiff2=1.0e-4 * dc**2 / 60.0, EddyDiff4=1.0e-4 * dc**4 / 600.0)
    kw.update(cfgkw)
    cfg = capi.TendConfig.default(**kw)
    h, u, tr = analytic.tc5(L, T)
    dt = 6.0 * max(1, int(0.25 * dc / 250.0 / 6.0))
    return L, cfg, (h, u, tr), dt


def plane_case(nx, K, T, **cfgkw):
    dc = 5.0e6 / 64
    m = plane_mesh(nx, nx, dc)
    L = decomp.build_all(m, 1, K)[0]
    kw = dict(ViscDel2=1.0e-4 * dc**2 / 60.0 * 10, ViscDel4=1.0e-3 * dc**4 / 600.0,
              EddyDiff2=1.0e-4 * dc**2 / 60.0, EddyDiff4=1.0e-4 * dc**4 / 600.0)
    kw.update(cfgkw)
    cfg = capi.TendConfig.default(**kw)
    h, u, tr = analytic.planar_state(L, T)
    dt = 6.0 * max(1, int(0.25 * dc / 100.0 / 6.0))
    return L, cfg, (h, u, tr), dt


@functools.lru_cache(maxsize=None)
def land_mesh(level):
    """The sphere with a continent (a cap of 25 degrees, which takes a pentagon with it), two islands of one cell and
    one of two cells removed: coastal edges with one cell, vertices with one or two cells, cells with one to three
    land neighbours."""
    m = sphere_mesh(level)
    d = np.arccos(np.clip(np.cos(m.latCell) * np.cos(0.3) * np.cos(m.lonCell - 1.0) + np.sin(m.latCell) * np.sin(0.3), -1, 1))
    keep = d > np.radians(25.0)
    far = np.nonzero(d > np.radians(60.0))[0]
    islands = [far[7], far[far.size // 2]]
    pair = far[far.size // 3]
    islands += [pair, m.cellsOnCell[pair, 0]]
    keep[islands] = False
    return meshgen.cull_cells(m, keep)


def land_case(level, K, T, **cfgkw):
    """TC5-style state on the mesh with land; the flow runs into the coast (u is not zero on coastal edges at t = 0,
    their tendency is masked), every Default.yml term on plus tracer hyperdiffusion."""
    m = land_mesh(level)
    L = decomp.build_all(m, 1, K)[0]
    dc = float(np.mean(m.dcEdge))
    kw = dict(ViscDel2=1.0e-4 * dc**2 / 60.0 * 10, ViscDel4=1.0e-3 * dc**4 / 600.0,
              EddyDiff2=1.0e-4 * dc**2 / 60.0, EddyDiff4=1.0e-4 * dc**4 / 600.0)
    kw.update(cfgkw)
    cfg = capi.TendConfig.default(**kw)
    h, u, tr = analytic.tc5(L, T)
    dt = 6.0 * max(1, int(0.25 * dc / 250.0 / 6.0))
    return L, cfg, (h, u, tr), dt


@functools.lru_cache(maxsize=None)
def hept_mesh(level):
    """12 heptagons and 12 extra pentagons (maxEdges = 7, up to 12 edgesOnEdge): meshgen.icosahedral_with_heptagons"""
    m = meshgen.icosahedral_with_heptagons(level, 6)
    m.bottomDepth = analytic.tc5_bottom_depth(m)
    return m


def hept_case(level, K, T, **cfgkw):
    m = hept_mesh(level)
    L = decomp.build_all(m, 1, K)[0]
    dc = float(np.mean(m.dcEdge))
    kw = dict(ViscDel2=1.0e-4 * dc**2 / 60.0 * 10, ViscDel4=1.0e-3 * dc**4 / 600.0,
              EddyDiff2=1.0e-4 * dc**2 / 60.0, EddyDiff4=1.0e-4 * dc**4 / 600.0)
    kw.update(cfgkw)
    cfg = capi.TendConfig.default(**kw)
    h, u, tr = analytic.tc5(L, T)
    dt = 6.0 * max(1, int(0.1 * dc / 250.0 / 6.0))       # the flipped pairs have short dual edges
    return L, cfg, (h, u, tr), dt


def all_terms_case(level=3, K=4, T=2, land=False):
    """Wind forcing, bottom drag, upwind fluxes on as well."""
    L, cfg, st, dt = (land_case if land else sphere_case)(level, K, T, WindForcingTendencyEnable=1, BottomDragTendencyEnable=1,
                                                          BottomDragCoeff=1.0e-3, FluxThicknessType=1, FluxTracerType=1)
    zonal = 0.1 * np.cos(L.LatCell)
    merid = 0.05 * np.sin(2 * L.LonCell) * np.cos(L.LatCell)
    zonal[-1] = merid[-1] = 0.0
    return L, cfg, st, dt, (np.ascontiguousarray(zonal), np.ascontiguousarray(merid))
