"""Analytic initial states for the synthetic meshes (BASELINE.json configs).

* ``tc5``  -- Williamson et al. (1992) test case 5 style: zonal flow over an
  isolated mountain (spec as in the reference tree's
  components/mpas-framework/src/core_sw/mpas_sw_test_cases.F:366), the
  mountain folded into bottomDepth so that Omega's per-layer
  ``SshCell = h - BottomDepth`` (LayerThicknessAuxVars.h:63-82) is the free
  surface.  Layers get slightly different velocities so a level-index bug
  cannot hide.
* ``tc2``  -- Williamson test case 2 (mpas_sw_test_cases.F:223): solid-body zonal flow in geostrophic balance over a
  flat bottom, an exact STEADY solution of the shallow-water equations, so whatever changes in a run is the
  discretisation error (tests/test_tc2_steady.py).
* ``manufactured`` -- initial condition of the manufactured shallow-water
  solution whose source terms are src/ocn/CustomTendencyTerms.cpp:107-210
  (doc/design/OmegaV0ShallowWater.md:656-666), config 1.
* tracers: ``2 - cos(lon + l) cos^4(lat)`` (variant of
  test/ocn/TendenciesTest.cpp:42-44), last tracer == 1 (conservation probe).
"""
from __future__ import annotations

import numpy as np

from .meshgen import OMEGA_EARTH

GRAV = 9.80665


def tc5_bottom_depth(mesh, h0=5960.0, hs0=2000.0):
    """bottomDepth = h0 - mountain; call before decomposition."""
    lon, lat = mesh.lonCell, mesh.latCell
    R = np.pi / 9.0
    dlon = np.abs(lon - 1.5 * np.pi)
    dlon = np.minimum(dlon, 2 * np.pi - dlon)
    r = np.minimum(R, np.sqrt(dlon**2 + (lat - np.pi / 6.0) ** 2))
    return h0 - hs0 * (1.0 - r / R)


TC2_U0 = 2.0 * np.pi * 6371220.0 / (12.0 * 86400.0)   # one revolution in 12 days
TC2_H0 = 2.94e4 / GRAV                                  # g h0 = 2.94e4 m^2 / s^2


def tc2_bottom_depth(mesh, h0=TC2_H0):
    """flat bottom at the reference depth (SshCell = h - BottomDepth is then the free-surface deviation)"""
    return np.full(mesh.nCells, h0)


def tc2(L, ntracers, u0=TC2_U0, h0=TC2_H0):
    """(h, u, tracers) of the steady zonal flow; every layer the same solution (no level factor: a layer with a
    different velocity would not be in balance with its own thickness)."""
    K = int(L.NVertLayers)
    if np.ptp(L.BottomDepth[:-1]) != 0.0:
        raise ValueError("tc2 needs a flat bottom: set mesh.bottomDepth = tc2_bottom_depth(mesh) before decomposing")
    a = float(L.sphere_radius)
    surf = h0 - (a * OMEGA_EARTH * u0 + 0.5 * u0 * u0) * np.sin(L.LatCell) ** 2 / GRAV
    depth = surf - (h0 - L.BottomDepth)
    h = np.repeat(depth[:, None], K, 1)
    u = np.repeat((u0 * np.cos(L.LatEdge) * np.cos(L.AngleEdge))[:, None], K, 1)
    h[-1], u[-1] = 0.0, 0.0
    _, _, tr = tc5(L, ntracers)
    return np.ascontiguousarray(h), np.ascontiguousarray(u), tr


def tc5(L, ntracers, u0=20.0, h0=5960.0):
    """(h, u, tracers) in host order for a local mesh L (sphere)."""
    K = int(L.NVertLayers)
    a = float(L.sphere_radius)
    latc, lonc = L.LatCell, L.LonCell
    late = L.LatEdge
    lev = 1.0 + 0.1 * np.arange(K) / max(K, 1)
    # surface height of the balanced zonal flow, minus the mountain = fluid depth
    surf = h0 - (a * OMEGA_EARTH * u0 + 0.5 * u0 * u0) * np.sin(latc) ** 2 / GRAV
    depth = surf - (h0 - L.BottomDepth)
    h = np.repeat(depth[:, None], K, 1) * (1.0 + 0.01 * (lev[None, :] - 1.0))
    un = u0 * np.cos(late) * np.cos(L.AngleEdge)
    u = un[:, None] * lev[None, :]
    h[-1], u[-1] = 0.0, 0.0
    tr = np.zeros((ntracers, latc.size, K))
    for l in range(ntracers):
        if l == ntracers - 1:
            tr[l] = 1.0
        else:
            tr[l] = (2.0 - np.cos(lonc + l) * np.cos(latc) ** 4)[:, None] * (1.0 + 0.05 * np.sin(0.3 * np.arange(K) + l))[None, :]
    if ntracers:
        tr[:, -1] = 0.0
    return np.ascontiguousarray(h), np.ascontiguousarray(u), np.ascontiguousarray(tr)


def planar_state(L, ntracers, h0=1000.0, amp=20.0, u0=1.0):
    """Smooth doubly periodic state for the planar hex meshes (all terms active)."""
    K = int(L.NVertLayers)
    Lx, Ly = float(L.Lx), float(L.Ly)
    kx, ky = 2 * np.pi / Lx, 2 * np.pi / Ly
    lev = 1.0 + 0.1 * np.arange(K) / max(K, 1)
    xc, yc, xe, ye = L.XCell, L.YCell, L.XEdge, L.YEdge
    h = (h0 + amp * np.sin(kx * xc) * np.cos(ky * yc))[:, None] * (1.0 + 0.01 * (lev - 1.0))[None, :]
    ux = u0 * np.sin(kx * xe) * np.cos(ky * ye)
    uy = u0 * np.cos(kx * xe) * np.sin(2 * ky * ye)
    u = (np.cos(L.AngleEdge) * ux + np.sin(L.AngleEdge) * uy)[:, None] * lev[None, :]
    h[-1], u[-1] = 0.0, 0.0
    tr = np.zeros((ntracers, xc.size, K))
    for l in range(ntracers):
        if l == ntracers - 1:
            tr[l] = 1.0
        else:
            tr[l] = (2.0 + np.cos(kx * xc + l) * np.sin(ky * yc))[:, None] * (1.0 + 0.05 * np.sin(0.3 * np.arange(K) + l))[None, :]
    if ntracers:
        tr[:, -1] = 0.0
    return np.ascontiguousarray(h), np.ascontiguousarray(u), np.ascontiguousarray(tr)


def manufactured(L, cfg, t=0.0):
    """h, u of the manufactured solution at time t (doc/design/OmegaV0ShallowWater.md:656-666):
    eta = eta0 sin(kx x + ky y - w t), u = v = eta0 cos(phase), h = H0 + eta."""
    K = int(L.NVertLayers)
    H0 = float(L.BottomDepth[0])
    kx, ky = 2 * np.pi / cfg.ManufacturedWavelengthX, 2 * np.pi / cfg.ManufacturedWavelengthY
    w = np.sqrt(H0 * GRAV * (kx * kx + ky * ky))
    eta0 = cfg.ManufacturedAmplitude
    ph_c = kx * L.XCell + ky * L.YCell - w * t
    ph_e = kx * L.XEdge + ky * L.YEdge - w * t
    h = np.repeat((H0 + eta0 * np.sin(ph_c))[:, None], K, 1)
    uv = eta0 * np.cos(ph_e)
    u = np.repeat((np.cos(L.AngleEdge) * uv + np.sin(L.AngleEdge) * uv)[:, None], K, 1)
    h[-1], u[-1] = 0.0, 0.0
    return np.ascontiguousarray(h), np.ascontiguousarray(u)
