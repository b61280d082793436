"""Python mirror of the Omega-1 column pieces (include/omega_b200.h, csrc/vertical.cu): VertCoord's four compute
methods, Eos::computeSpecVol / computeSpecVolDisp and the batched tridiagonal solvers, with the reference's
method names.  Each method forwards to one C entry point; torch supplies device memory only."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import capi
from .model import HorzMesh, _chk, _dptr, _stream

LINEAR_EOS, TEOS10_EOS = 0, 1


def _iptr(t):
    return C.cast(C.c_void_p(t.data_ptr()), capi.c_int_p)


class VertCoord:
    """src/ocn/VertCoord.h: active layer range per column + the diagnostic column integrals."""

    def __init__(self, mesh: HorzMesh, min_layer_cell: np.ndarray, max_layer_cell: np.ndarray, rho0: float = 1026.0):
        self.mesh, self.lib, self.Rho0 = mesh, mesh.lib, float(rho0)
        p, _ = mesh.perm(capi.ON_CELL)
        K, n = mesh.K, mesh.NSize[0]
        self.MinLayerCell = torch.from_numpy(np.ascontiguousarray(np.asarray(min_layer_cell, np.int32)[p])).to(mesh.device)
        self.MaxLayerCell = torch.from_numpy(np.ascontiguousarray(np.asarray(max_layer_cell, np.int32)[p])).to(mesh.device)
        z = lambda w: torch.zeros((n, w), dtype=mesh.dtype, device=mesh.device)
        self.PressureInterface, self.PressureMid = z(K + 1), z(K)
        self.ZInterface, self.ZMid, self.GeopotentialMid, self.LayerThicknessTarget = z(K + 1), z(K), z(K), z(K)

    def computePressure(self, layer_thickness, surface_pressure):
        _chk(self.lib, self.lib.omega_vert_compute_pressure(self.mesh.h, _iptr(self.MinLayerCell), _iptr(self.MaxLayerCell),
             C.c_double(self.Rho0), _dptr(layer_thickness), _dptr(surface_pressure), _dptr(self.PressureInterface),
             _dptr(self.PressureMid), _stream()), "omega_vert_compute_pressure")

    def computeZHeight(self, layer_thickness, spec_vol):
        _chk(self.lib, self.lib.omega_vert_compute_z_height(self.mesh.h, _iptr(self.MinLayerCell), _iptr(self.MaxLayerCell),
             C.c_double(self.Rho0), _dptr(layer_thickness), _dptr(spec_vol), _dptr(self.ZInterface), _dptr(self.ZMid),
             _stream()), "omega_vert_compute_z_height")

    def computeGeopotential(self, tidal_potential, self_attraction_loading):
        _chk(self.lib, self.lib.omega_vert_compute_geopotential(self.mesh.h, _iptr(self.MinLayerCell), _iptr(self.MaxLayerCell),
             _dptr(self.ZMid), _dptr(tidal_potential), _dptr(self_attraction_loading), _dptr(self.GeopotentialMid),
             _stream()), "omega_vert_compute_geopotential")

    def computeTargetThickness(self, ref_layer_thickness, movement_weights):
        _chk(self.lib, self.lib.omega_vert_compute_target_thickness(self.mesh.h, _iptr(self.MinLayerCell), _iptr(self.MaxLayerCell),
             C.c_double(self.Rho0), _dptr(self.PressureInterface), _dptr(ref_layer_thickness), _dptr(movement_weights),
             _dptr(self.LayerThicknessTarget), _stream()), "omega_vert_compute_target_thickness")


class Eos:
    """src/ocn/Eos.h: specific volume, linear or TEOS-10 (Roquet et al. 2015)."""

    def __init__(self, mesh: HorzMesh, eos_choice: int = TEOS10_EOS, drho_dt=-0.2, drho_ds=0.8, rho_t0s0=1000.0):
        self.mesh, self.lib, self.EosChoice = mesh, mesh.lib, eos_choice
        self.DRhodT, self.DRhodS, self.RhoT0S0 = drho_dt, drho_ds, rho_t0s0
        self.SpecVol = mesh.alloc(capi.ON_CELL)
        self.SpecVolDisplaced = mesh.alloc(capi.ON_CELL)

    def _run(self, out, ct, sa, p, kdisp):
        _chk(self.lib, self.lib.omega_eos_compute_spec_vol(self.mesh.h, self.EosChoice, _dptr(ct), _dptr(sa),
             _dptr(p) if p is not None else None, kdisp, C.c_double(self.DRhodT), C.c_double(self.DRhodS),
             C.c_double(self.RhoT0S0), _dptr(out), _stream()), "omega_eos_compute_spec_vol")

    def computeSpecVol(self, conserv_temp, abs_salinity, pressure):
        self._run(self.SpecVol, conserv_temp, abs_salinity, pressure, 0)

    def computeSpecVolDisp(self, conserv_temp, abs_salinity, pressure, kdisp):
        self._run(self.SpecVolDisplaced, conserv_temp, abs_salinity, pressure, kdisp)


def tridiag_solve(lib, dl, d, du, x):
    """TriDiagSolver::solve: [NBatch][NRow] device tensors, x = rhs on entry, solution on return."""
    nb, nr = x.shape
    _chk(lib, lib.omega_tridiag_solve(nb, nr, _dptr(dl), _dptr(d), _dptr(du), _dptr(x), _stream()), "omega_tridiag_solve")


def tridiag_diffusion_solve(lib, g, h, x):
    """TriDiagDiffSolver::solve: D = H + G(k-1) + G(k), off-diagonals -G."""
    nb, nr = x.shape
    _chk(lib, lib.omega_tridiag_diffusion_solve(nb, nr, _dptr(g), _dptr(h), _dptr(x), _stream()), "omega_tridiag_diffusion_solve")
