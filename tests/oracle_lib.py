"""ctypes wrapper of the CPU oracle (oracle/libomega_oracle*.so, and
oracle/_ref/libomega_ref.so which exports the same symbols).  TEST
INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from omega_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
STRICT = os.path.join(ORACLE_DIR, "libomega_oracle.so")
FAST = os.path.join(ORACLE_DIR, "libomega_oracle_fast.so")
REF = os.path.join(ORACLE_DIR, "_ref", "libomega_ref.so")
REF_FAST = os.path.join(ORACLE_DIR, "_ref", "libomega_ref_fast.so")  # -O3, OMEGA_VECTOR_LENGTH=8


def build_oracle():
    subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)


_dp = capi.c_dbl_p


def _p(a):
    assert a.dtype == np.float64 and a.flags.c_contiguous
    return a.ctypes.data_as(_dp)


class Oracle:
    """One OracleStepper over a decomp.build_local() mesh (host order)."""

    def __init__(self, L, cfg: capi.TendConfig, ntracers: int, lib_path: str = STRICT):
        if not os.path.exists(lib_path):
            build_oracle()
        self.lib = C.CDLL(lib_path)
        self.L, self.cfg, self.T = L, cfg, ntracers
        self.K = int(L.NVertLayers)
        self.desc, self._keep = capi.make_mesh_desc(L)
        self.h = C.c_void_p()
        rc = self.lib.oracle_stepper_create(C.byref(self.desc), C.byref(cfg), ntracers, C.byref(self.h))
        assert rc == 0
        aux = capi.AuxArrays()
        self.lib.oracle_stepper_aux_arrays(self.h, C.byref(aux))
        tend = capi.TendArrays()
        self.lib.oracle_stepper_tend_arrays(self.h, C.byref(tend))
        self.aux = {n: self._view(getattr(aux, n), *capi.AUX_KIND[n]) for n in capi.AUX_FIELDS}
        self.tend = {"LayerThicknessTend": self._view(tend.LayerThicknessTend, capi.ON_CELL, 1, 0),
                     "NormalVelocityTend": self._view(tend.NormalVelocityTend, capi.ON_EDGE, 1, 0),
                     "TracerTend": self._view(tend.TracerTend, capi.ON_CELL, 1, 1)}

    def nsize(self, kind):
        return int((self.L.NCellsSize, self.L.NEdgesSize, self.L.NVerticesSize)[kind])

    def _view(self, ptr, kind, has_k, has_t):
        shape = ((self.T,) if has_t else ()) + (self.nsize(kind),) + ((self.K,) if has_k else ())
        n = int(np.prod(shape))
        if n == 0:
            return np.zeros(shape)
        return np.ctypeslib.as_array(ptr, shape=(n,)).reshape(shape)

    def set_custom_tendencies(self, thickness=None, velocity=None):
        """fn(tend, h, u, time) on numpy views (host order; h is None in the ...Only evaluations of the C oracle) that
        ADD into tend: Tendencies::CustomThicknessTend / CustomVelocityTend (Tendencies.h:51-53)."""
        def make(fn, kind):
            if fn is None:
                return capi.CUSTOM_TENDENCY(0)

            def cb(user, tend, h, u, time, stream):
                dp = C.POINTER(C.c_double)
                v = lambda p, k: None if not p else self._view(C.cast(p, dp), k, 1, 0)
                fn(v(tend, kind), v(h, capi.ON_CELL), v(u, capi.ON_EDGE), float(time))
                return 0
            return capi.CUSTOM_TENDENCY(cb)
        self._custom = (make(thickness, capi.ON_CELL), make(velocity, capi.ON_EDGE))
        rc = self.lib.oracle_stepper_set_custom_tendencies(self.h, self._custom[0], self._custom[1], None)
        assert rc == 0, rc
        tend = capi.TendArrays()
        self.lib.oracle_stepper_tend_arrays(self.h, C.byref(tend))     # the reference build switches to another object
        self.tend = {"LayerThicknessTend": self._view(tend.LayerThicknessTend, capi.ON_CELL, 1, 0),
                     "NormalVelocityTend": self._view(tend.NormalVelocityTend, capi.ON_EDGE, 1, 0),
                     "TracerTend": self._view(tend.TracerTend, capi.ON_CELL, 1, 1)}

    def close(self):
        if self.h:
            self.lib.oracle_stepper_destroy(self.h)
            self.h = None

    def set_config(self, cfg):
        self.cfg = cfg
        self.lib.oracle_stepper_set_config(self.h, C.byref(cfg))

    def derived(self):
        esc = np.zeros((self.L.NCellsSize, self.L.MaxEdges))
        esv = np.zeros((self.L.NVerticesSize, self.L.VertexDegree))
        mask = np.zeros(self.L.NEdgesSize)
        self.lib.oracle_mesh_derived(self.h, _p(esc), _p(esv), _p(mask))
        return esc, esv, mask

    def set_wind_stress(self, zonal, merid):
        self.lib.oracle_stepper_set_wind_stress(self.h, _p(zonal), _p(merid))

    def aux_compute_mom(self, h, u):
        self.lib.oracle_aux_compute_mom(self.h, _p(h), _p(u))

    def aux_compute_all(self, h, u, tr):
        self.lib.oracle_aux_compute_all(self.h, _p(h), _p(u), _p(tr))

    def tend_all(self, h, u, tr, time=0.0, u_raw=None):
        self.lib.oracle_tend_all(self.h, _p(h), _p(u), _p(u if u_raw is None else u_raw), _p(tr),
                                 C.c_double(time))

    def tend_thickness_only(self, u, time=0.0):
        self.lib.oracle_tend_thickness_only(self.h, _p(u), C.c_double(time))

    def tend_velocity_only(self, u, time=0.0, u_raw=None):
        self.lib.oracle_tend_velocity_only(self.h, _p(u), _p(u if u_raw is None else u_raw), C.c_double(time))

    def tend_tracer_only(self, u_raw, tr):
        self.lib.oracle_tend_tracer_only(self.h, _p(u_raw), _p(tr))

    def make_state(self, h0, u0, tr0):
        """StateArrays over fresh numpy ring buffers (slot 0 = current)."""
        self.hbuf = [np.array(h0), np.zeros_like(h0)]
        self.ubuf = [np.array(u0), np.zeros_like(u0)]
        self.tbuf = [np.array(tr0), np.zeros_like(tr0)]
        st = capi.StateArrays(NTimeLevels=2, CurTimeIndex=0, TracerCurTimeIndex=0, NTracers=self.T)
        for i in range(2):
            st.LayerThickness[i] = _p(self.hbuf[i])
            st.NormalVelocity[i] = _p(self.ubuf[i])
            st.Tracers[i] = _p(self.tbuf[i]) if self.T else None
        self.state = st
        return st

    def do_step(self, stepper, coeffs, compat=1):
        self.lib.oracle_stepper_do_step(self.h, stepper, C.byref(self.state), C.byref(coeffs), 0, compat)

    def current(self):
        i, j = self.state.CurTimeIndex, self.state.TracerCurTimeIndex
        return self.hbuf[i], self.ubuf[i], self.tbuf[j]

    # generic operators for the known-answer tests
    def op(self, name, src, dst, *extra):
        getattr(self.lib, "oracle_op_" + name)(self.h, _p(src), *extra, _p(dst))

    # ---- phase-wise stepping for the multi-rank emulation (C oracle only) ----
    def step_phase(self, stepper, coeffs, phase, compat=1):
        self.lib.oracle_stepper_step_phase(self.h, stepper, C.byref(self.state), C.byref(coeffs), phase, 0, compat)

    def exchange_arrays(self, stepper, phase):
        """numpy views (host order) of the arrays to halo-exchange after `phase`:
        [(kind name, array)]"""
        ptrs = (_dp * 3)()
        elems, nouter, n = (C.c_int * 3)(), (C.c_int * 3)(), C.c_int()
        self.lib.oracle_stepper_exchange_list(self.h, stepper, C.byref(self.state), phase, ptrs, elems, nouter, C.byref(n))
        out = []
        for i in range(n.value):
            kind = int(elems[i])
            has_t = nouter[i] > 1 or (i == 2)
            if i == 2 and self.T == 0:
                continue
            out.append((("Cell", "Edge", "Vertex")[kind], self._view(ptrs[i], kind, 1, 1 if i == 2 else 0)))
        return out

    def rotate(self):
        self.lib.oracle_state_rotate(C.byref(self.state))
