"""Python host mirror over the C ABI (include/omega_b200.h).

Mirrors the reference's object model for the hot path -- HorzMesh,
OceanState/Tracers time-level rings, AuxiliaryState, Tendencies, TimeStepper
-- with the same method names; every method forwards to one C entry point of
libomega_b200.so.  torch supplies device memory and streams only.  There is
no CPU fallback: construction raises when the CUDA library or a GPU is
missing.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import capi
from .capi import ON_CELL, ON_EDGE, ON_VERTEX


def _chk(lib, rc, what):
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {lib.omega_last_error().decode()}")


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dptr(t: torch.Tensor):
    return C.cast(C.c_void_p(t.data_ptr()), capi.c_dbl_p)


class HorzMesh:
    """Device-resident mesh (omega_mesh_create); src/ocn/HorzMesh.h."""

    def __init__(self, L, reorder: bool = True, device: str | torch.device = "cuda:0", lib=None):
        if not torch.cuda.is_available():
            raise RuntimeError("omega_b200 needs a CUDA device; there is no CPU path")
        self.lib = lib or capi.load_library()
        # Real of the loaded build: fp64, or fp32 for libomega_b200_sp.so (capi.load_library(capi.LIB_PATH_SP))
        self.real_bytes = int(self.lib.omega_b200_real_bytes())
        self.dtype = torch.float64 if self.real_bytes == 8 else torch.float32
        self.np_dtype = np.float64 if self.real_bytes == 8 else np.float32
        self.device = torch.device(device)
        torch.cuda.set_device(self.device)
        self.L = L
        self.K = int(L.NVertLayers)
        self.desc, self._keep = capi.make_mesh_desc(L)
        self.h = C.c_void_p()
        _chk(self.lib, self.lib.omega_mesh_create(C.byref(self.desc), int(reorder), C.byref(self.h)), "omega_mesh_create")
        self.NSize = (int(L.NCellsSize), int(L.NEdgesSize), int(L.NVerticesSize))
        self.NAll = (int(L.NCellsAll), int(L.NEdgesAll), int(L.NVerticesAll))
        self.NOwned = (int(L.NCellsOwned), int(L.NEdgesOwned), int(L.NVerticesOwned))

    def perm(self, kind):
        p = np.zeros(self.NSize[kind], np.int32)
        ip = np.zeros(self.NSize[kind], np.int32)
        _chk(self.lib, self.lib.omega_mesh_perm(self.h, kind, p.ctypes.data_as(capi.c_int_p), ip.ctypes.data_as(capi.c_int_p)), "omega_mesh_perm")
        return p, ip

    def derived(self):
        esc = np.zeros((self.NSize[0], int(self.L.MaxEdges)))
        esv = np.zeros((self.NSize[2], 3))
        mask = np.zeros(self.NSize[1])
        f = lambda a: a.ctypes.data_as(capi.c_dbl_p)
        _chk(self.lib, self.lib.omega_mesh_derived(self.h, f(esc), f(esv), f(mask)), "omega_mesh_derived")
        return esc, esv, mask

    def alloc(self, kind, nouter=0, levels=True):
        shape = ((nouter,) if nouter else ()) + (self.NSize[kind],) + ((self.K,) if levels else ())
        return torch.zeros(shape, dtype=self.dtype, device=self.device)

    def to_device(self, kind, host: np.ndarray, out: torch.Tensor | None = None, sync: bool = True):
        """host-order numpy [..., NSize, K] -> device-order tensor.  sync=False: asynchronous on the
        current stream (host must be pinned and stay alive until the stream has passed the copy)."""
        host = np.ascontiguousarray(host, self.np_dtype)
        nouter = host.shape[0] if host.ndim == 3 else 1
        ninner = host.shape[-1] if host.ndim >= 2 else 1
        if out is None:
            out = torch.empty(host.shape, dtype=self.dtype, device=self.device)
        _chk(self.lib, self.lib.omega_field_to_device(self.h, kind, nouter, ninner, host.ctypes.data_as(capi.c_dbl_p), _dptr(out), _stream()), "omega_field_to_device")
        if sync:
            torch.cuda.current_stream().synchronize()  # host buffer may be a temporary
        return out

    def to_host(self, kind, dev: torch.Tensor, out: np.ndarray | None = None, sync: bool = True):
        """sync=False: asynchronous on the current stream into a pinned `out`."""
        nouter = dev.shape[0] if dev.dim() == 3 else 1
        ninner = dev.shape[-1] if dev.dim() >= 2 else 1
        if out is None:
            out = np.empty(tuple(dev.shape), self.np_dtype)
        fn = self.lib.omega_field_to_host if sync else self.lib.omega_field_to_host_async
        _chk(self.lib, fn(self.h, kind, nouter, ninner, _dptr(dev), out.ctypes.data_as(capi.c_dbl_p), _stream()), "omega_field_to_host")
        return out

    def close(self):
        if self.h:
            self.lib.omega_mesh_destroy(self.h)
            self.h = None


class OceanState:
    """h/u/tracer time-level rings on device (src/ocn/OceanState.cpp:65-117,
    348-365, 394-407; src/ocn/Tracers.cpp:269-277, 457-496)."""

    def __init__(self, mesh: HorzMesh, ntracers: int):
        self.mesh, self.T = mesh, ntracers
        self.LayerThickness = [mesh.alloc(ON_CELL) for _ in range(2)]
        self.NormalVelocity = [mesh.alloc(ON_EDGE) for _ in range(2)]
        self.Tracers = [mesh.alloc(ON_CELL, ntracers) if ntracers else None for _ in range(2)]
        self.c = capi.StateArrays(NTimeLevels=2, CurTimeIndex=0, TracerCurTimeIndex=0, NTracers=ntracers)
        for i in range(2):
            self.c.LayerThickness[i] = _dptr(self.LayerThickness[i])
            self.c.NormalVelocity[i] = _dptr(self.NormalVelocity[i])
            self.c.Tracers[i] = _dptr(self.Tracers[i]) if ntracers else None

    def time_index(self, level):
        return (level + self.c.CurTimeIndex + 2) % 2

    def tracer_time_index(self, level):
        return (level + self.c.TracerCurTimeIndex + 2) % 2

    def getLayerThickness(self, level):
        return self.LayerThickness[self.time_index(level)]

    def getNormalVelocity(self, level):
        return self.NormalVelocity[self.time_index(level)]

    def getTracers(self, level):
        return self.Tracers[self.tracer_time_index(level)]

    def copyToDevice(self, h, u, tr=None, level=0):
        self.mesh.to_device(ON_CELL, h, self.getLayerThickness(level))
        self.mesh.to_device(ON_EDGE, u, self.getNormalVelocity(level))
        if self.T:
            self.mesh.to_device(ON_CELL, tr, self.getTracers(level))

    def copyToHost(self, level=0):
        h = self.mesh.to_host(ON_CELL, self.getLayerThickness(level))
        u = self.mesh.to_host(ON_EDGE, self.getNormalVelocity(level))
        tr = self.mesh.to_host(ON_CELL, self.getTracers(level)) if self.T else None
        return h, u, tr


class Stepper:
    """AuxiliaryState + Tendencies + TimeStepper (omega_stepper_*)."""

    def __init__(self, mesh: HorzMesh, cfg: capi.TendConfig, ntracers: int):
        self.mesh, self.lib, self.T, self.cfg = mesh, mesh.lib, ntracers, cfg
        self.h = C.c_void_p()
        _chk(self.lib, self.lib.omega_stepper_create(mesh.h, C.byref(cfg), ntracers, C.byref(self.h)), "omega_stepper_create")
        self.lib.omega_stepper_launch_count.restype = C.c_longlong
        self._aux = None

    def set_config(self, cfg):
        self.cfg = cfg
        _chk(self.lib, self.lib.omega_stepper_set_config(self.h, C.byref(cfg)), "omega_stepper_set_config")

    # -- array access ---------------------------------------------------------
    def _wrap(self, ptr, kind, has_k, has_t):
        """numpy host-order copy of a library-owned device array."""
        m = self.mesh
        shape = ((self.T,) if has_t else ()) + (m.NSize[kind],) + ((m.K,) if has_k else ())
        n = int(np.prod(shape))
        if n == 0 or not ptr:
            return np.zeros(shape, m.np_dtype)
        out = np.empty(shape, m.np_dtype)
        nouter = self.T if has_t else 1
        ninner = m.K if has_k else 1
        _chk(self.lib, self.lib.omega_field_to_host(m.h, kind, nouter, ninner, ptr, out.ctypes.data_as(capi.c_dbl_p), _stream()), "omega_field_to_host")
        return out

    def aux_to_host(self, names=None):
        aux = capi.AuxArrays()
        _chk(self.lib, self.lib.omega_stepper_aux_arrays(self.h, C.byref(aux)), "omega_stepper_aux_arrays")
        return {n: self._wrap(getattr(aux, n), *capi.AUX_KIND[n]) for n in (names or capi.AUX_FIELDS)}

    def aux_from_host(self, name, host):
        aux = capi.AuxArrays()
        _chk(self.lib, self.lib.omega_stepper_aux_arrays(self.h, C.byref(aux)), "omega_stepper_aux_arrays")
        kind, has_k, has_t = capi.AUX_KIND[name]
        host = np.ascontiguousarray(host, self.mesh.np_dtype)
        _chk(self.lib, self.lib.omega_field_to_device(self.mesh.h, kind, self.T if has_t else 1, self.mesh.K if has_k else 1,
                                                      host.ctypes.data_as(capi.c_dbl_p), getattr(aux, name), _stream()), "omega_field_to_device")
        torch.cuda.current_stream().synchronize()

    def tend_to_host(self):
        t = capi.TendArrays()
        _chk(self.lib, self.lib.omega_stepper_tend_arrays(self.h, C.byref(t)), "omega_stepper_tend_arrays")
        return {"LayerThicknessTend": self._wrap(t.LayerThicknessTend, ON_CELL, 1, 0),
                "NormalVelocityTend": self._wrap(t.NormalVelocityTend, ON_EDGE, 1, 0),
                "TracerTend": self._wrap(t.TracerTend, ON_CELL, 1, 1)}

    def set_wind_stress(self, zonal, merid):
        z, m = np.ascontiguousarray(zonal, self.mesh.np_dtype), np.ascontiguousarray(merid, self.mesh.np_dtype)
        _chk(self.lib, self.lib.omega_stepper_set_wind_stress(self.h, z.ctypes.data_as(capi.c_dbl_p), m.ctypes.data_as(capi.c_dbl_p), _stream()), "set_wind_stress")
        torch.cuda.current_stream().synchronize()

    # -- custom tendencies (Tendencies.h:51-53) ------------------------------------
    def _device_tensor(self, ptr, kind):
        """torch view (no copy) of a library-owned [N*Size][K] device array in device order"""
        if not ptr:
            return None

        class _Dev:
            pass
        d = _Dev()
        m = self.mesh
        d.__cuda_array_interface__ = dict(shape=(m.NSize[kind], m.K), typestr="<f8" if m.real_bytes == 8 else "<f4",
                                          data=(int(ptr), False), version=2)
        return torch.as_tensor(d, device=m.device)

    def set_custom_tendencies(self, thickness=None, velocity=None):
        """thickness / velocity: callables fn(tend, h, u, time) on torch views of the device arrays (device order;
        h is None in the ...Only evaluations) that ADD their term into tend on torch's current stream -- the role of
        Tendencies::CustomThicknessTend / CustomVelocityTend.  None, None removes them."""
        def make(fn, kind):
            if fn is None:
                return capi.CUSTOM_TENDENCY(0)

            def cb(user, tend, h, u, time, stream):
                try:
                    fn(self._device_tensor(tend, kind), self._device_tensor(h, ON_CELL), self._device_tensor(u, ON_EDGE), float(time))
                    return 0
                except Exception as e:   # an exception must not cross the C frame
                    self._custom_error = e
                    return 1
            return capi.CUSTOM_TENDENCY(cb)
        self._custom_error = None
        self._custom = (make(thickness, ON_CELL), make(velocity, ON_EDGE))   # keep the trampolines alive
        _chk(self.lib, self.lib.omega_stepper_set_custom_tendencies(self.h, self._custom[0], self._custom[1], None),
             "omega_stepper_set_custom_tendencies")

    # -- AuxiliaryState / Tendencies ---------------------------------------------
    def computeMomAux(self, h, u):
        _chk(self.lib, self.lib.omega_aux_compute_mom(self.h, _dptr(h), _dptr(u), _stream()), "omega_aux_compute_mom")

    def computeAll(self, h, u, tr):
        _chk(self.lib, self.lib.omega_aux_compute_all(self.h, _dptr(h), _dptr(u), _dptr(tr) if tr is not None else None, _stream()), "omega_aux_compute_all")

    def computeAllTendencies(self, h, u, tr, time=0.0, u_raw=None):
        _chk(self.lib, self.lib.omega_tend_all(self.h, _dptr(h), _dptr(u), _dptr(u if u_raw is None else u_raw),
                                               _dptr(tr) if tr is not None else None, C.c_double(time), _stream()), "omega_tend_all")

    def computeThicknessTendenciesOnly(self, u, time=0.0):
        _chk(self.lib, self.lib.omega_tend_thickness_only(self.h, _dptr(u), C.c_double(time), _stream()), "omega_tend_thickness_only")

    def computeVelocityTendenciesOnly(self, u, time=0.0, u_raw=None):
        _chk(self.lib, self.lib.omega_tend_velocity_only(self.h, _dptr(u), _dptr(u if u_raw is None else u_raw), C.c_double(time), _stream()), "omega_tend_velocity_only")

    def computeTracerTendenciesOnly(self, u_raw, tr):
        _chk(self.lib, self.lib.omega_tend_tracer_only(self.h, _dptr(u_raw), _dptr(tr), _stream()), "omega_tend_tracer_only")

    # -- TimeStepper ----------------------------------------------------------------
    def doStep(self, state: OceanState, stepper: int, coeffs: capi.StepCoeffs, fused=True, compat=True):
        _chk(self.lib, self.lib.omega_stepper_do_step(self.h, stepper, C.byref(state.c), C.byref(coeffs), None,
                                                      int(fused), int(compat), _stream()), "omega_stepper_do_step")

    def stepPhase(self, state, stepper, coeffs, phase, fused=True, compat=True):
        _chk(self.lib, self.lib.omega_stepper_step_phase(self.h, stepper, C.byref(state.c), C.byref(coeffs), phase,
                                                         int(fused), int(compat), _stream()), "omega_stepper_step_phase")

    def launch_count(self):
        return int(self.lib.omega_stepper_launch_count(self.h))

    def last_path(self):
        """OMEGA_PATH_* bits of the last step: 1 fused, 2 tiled cell pass, 4 k_fCT2, 8 tiled edge pass."""
        return int(self.lib.omega_stepper_last_path(self.h))

    def close(self):
        if self.h:
            self.lib.omega_stepper_destroy(self.h)
            self.h = None
