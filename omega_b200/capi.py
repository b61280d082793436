"""ctypes view of include/omega_b200.h (structs shared by the GPU library and,
in tests only, by the CPU oracle which reuses the same struct layouts)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

c_int_p = C.POINTER(C.c_int)
c_dbl_p = C.POINTER(C.c_double)

ON_CELL, ON_EDGE, ON_VERTEX = 0, 1, 2
# OmegaCustomTendency: int (*)(void *User, Real *Tend, const Real *H, const Real *U, double Time, void *Stream)
CUSTOM_TENDENCY = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p)
RK4, RK2, FB = 0, 1, 2


class MeshDesc(C.Structure):
    _fields_ = (
        [(n, C.c_int) for n in (
            "NCellsOwned", "NCellsAll", "NCellsSize", "NEdgesOwned", "NEdgesAll", "NEdgesSize",
            "NVerticesOwned", "NVerticesAll", "NVerticesSize", "MaxEdges", "MaxEdges2",
            "VertexDegree", "NVertLayers")]
        + [(n, c_int_p) for n in (
            "NEdgesOnCell", "EdgesOnCell", "CellsOnEdge", "VerticesOnEdge", "NEdgesOnEdge",
            "EdgesOnEdge", "CellsOnVertex", "EdgesOnVertex")]
        + [(n, c_dbl_p) for n in (
            "AreaCell", "BottomDepth", "XCell", "YCell", "ZCell",
            "DcEdge", "DvEdge", "AngleEdge", "FEdge", "XEdge", "YEdge",
            "WeightsOnEdge", "AreaTriangle", "FVertex", "KiteAreasOnVertex")])


_CFG_INTS = ("ThicknessFluxTendencyEnable", "PVTendencyEnable", "KETendencyEnable",
             "SSHTendencyEnable", "VelDiffTendencyEnable", "VelHyperDiffTendencyEnable",
             "WindForcingTendencyEnable", "BottomDragTendencyEnable",
             "TracerHorzAdvTendencyEnable", "TracerDiffTendencyEnable",
             "TracerHyperDiffTendencyEnable", "FluxThicknessType", "FluxTracerType",
             "WindStressInterpType", "ManufacturedSolutionTendency")
_CFG_DBLS = ("ViscDel2", "ViscDel4", "DivFactor", "Density0", "BottomDragCoeff", "EddyDiff2",
             "EddyDiff4", "ManufacturedWavelengthX", "ManufacturedWavelengthY",
             "ManufacturedAmplitude")


class TendConfig(C.Structure):
    _fields_ = [(n, C.c_int) for n in _CFG_INTS] + [(n, C.c_double) for n in _CFG_DBLS]

    @classmethod
    def default(cls, **kw):
        """configs/Default.yml:32-52 values."""
        c = cls(ThicknessFluxTendencyEnable=1, PVTendencyEnable=1, KETendencyEnable=1,
                SSHTendencyEnable=1, VelDiffTendencyEnable=1, VelHyperDiffTendencyEnable=1,
                WindForcingTendencyEnable=0, BottomDragTendencyEnable=0,
                TracerHorzAdvTendencyEnable=1, TracerDiffTendencyEnable=1,
                TracerHyperDiffTendencyEnable=1, FluxThicknessType=0, FluxTracerType=0,
                WindStressInterpType=0, ManufacturedSolutionTendency=0,
                ViscDel2=1.0e3, ViscDel4=1.2e11, DivFactor=1.0, Density0=1026.0,
                BottomDragCoeff=0.0, EddyDiff2=10.0, EddyDiff4=0.0,
                ManufacturedWavelengthX=5.0e6, ManufacturedWavelengthY=4.33013e6,
                ManufacturedAmplitude=1.0)
        for k, v in kw.items():
            if not hasattr(c, k):
                raise AttributeError(k)
            setattr(c, k, v)
        return c


AUX_FIELDS = ("KineticEnergyCell", "VelocityDivCell", "FluxLayerThickEdge", "MeanLayerThickEdge",
              "SshCell", "RelVortVertex", "NormRelVortVertex", "NormPlanetVortVertex",
              "NormRelVortEdge", "NormPlanetVortEdge", "Del2Edge", "Del2DivCell",
              "Del2RelVortVertex", "HTracersEdge", "Del2TracersCell", "NormalStressEdge",
              "ZonalStressCell", "MeridStressCell")
# element kind, has level dim, has tracer dim
AUX_KIND = {"KineticEnergyCell": (ON_CELL, 1, 0), "VelocityDivCell": (ON_CELL, 1, 0),
            "FluxLayerThickEdge": (ON_EDGE, 1, 0), "MeanLayerThickEdge": (ON_EDGE, 1, 0),
            "SshCell": (ON_CELL, 1, 0), "RelVortVertex": (ON_VERTEX, 1, 0),
            "NormRelVortVertex": (ON_VERTEX, 1, 0), "NormPlanetVortVertex": (ON_VERTEX, 1, 0),
            "NormRelVortEdge": (ON_EDGE, 1, 0), "NormPlanetVortEdge": (ON_EDGE, 1, 0),
            "Del2Edge": (ON_EDGE, 1, 0), "Del2DivCell": (ON_CELL, 1, 0),
            "Del2RelVortVertex": (ON_VERTEX, 1, 0), "HTracersEdge": (ON_EDGE, 1, 1),
            "Del2TracersCell": (ON_CELL, 1, 1), "NormalStressEdge": (ON_EDGE, 0, 0),
            "ZonalStressCell": (ON_CELL, 0, 0), "MeridStressCell": (ON_CELL, 0, 0)}


class AuxArrays(C.Structure):
    _fields_ = [(n, c_dbl_p) for n in AUX_FIELDS]


class TendArrays(C.Structure):
    _fields_ = [(n, c_dbl_p) for n in ("LayerThicknessTend", "NormalVelocityTend", "TracerTend")]


class StateArrays(C.Structure):
    _fields_ = [("NTimeLevels", C.c_int), ("CurTimeIndex", C.c_int),
                ("TracerCurTimeIndex", C.c_int), ("NTracers", C.c_int),
                ("LayerThickness", c_dbl_p * 2), ("NormalVelocity", c_dbl_p * 2),
                ("Tracers", c_dbl_p * 2)]


class StepCoeffs(C.Structure):
    _fields_ = [("NStages", C.c_int), ("A", C.c_double * 4), ("B", C.c_double * 4),
                ("StageTime", C.c_double * 4), ("Dt", C.c_double)]


class HaloDesc(C.Structure):
    _fields_ = [("MyTask", C.c_int), ("NNeighbors", C.c_int), ("NeighborTasks", c_int_p),
                ("SendOffsets", c_int_p * 3), ("SendIdx", c_int_p * 3),
                ("RecvOffsets", c_int_p * 3), ("RecvIdx", c_int_p * 3)]


def _ip(a):
    return a.ctypes.data_as(c_int_p)


def _dp(a):
    return a.ctypes.data_as(c_dbl_p)


def make_mesh_desc(L):
    """MeshDesc over the numpy arrays of a decomp.build_local() mesh.
    Returns (desc, keepalive)."""
    d = MeshDesc()
    keep = []
    for n in ("NCellsOwned", "NCellsAll", "NCellsSize", "NEdgesOwned", "NEdgesAll", "NEdgesSize",
              "NVerticesOwned", "NVerticesAll", "NVerticesSize", "MaxEdges", "MaxEdges2",
              "VertexDegree", "NVertLayers"):
        setattr(d, n, int(L[n]))
    for n in ("NEdgesOnCell", "EdgesOnCell", "CellsOnEdge", "VerticesOnEdge", "NEdgesOnEdge",
              "EdgesOnEdge", "CellsOnVertex", "EdgesOnVertex"):
        a = np.ascontiguousarray(L[n], np.int32)
        keep.append(a)
        setattr(d, n, _ip(a))
    for n in ("AreaCell", "BottomDepth", "XCell", "YCell", "ZCell", "DcEdge", "DvEdge",
              "AngleEdge", "FEdge", "XEdge", "YEdge", "WeightsOnEdge", "AreaTriangle", "FVertex",
              "KiteAreasOnVertex"):
        a = np.ascontiguousarray(L[n], np.float64)
        keep.append(a)
        setattr(d, n, _dp(a))
    return d, keep


def rk_coeffs(stepper: int, dt: float, t0: float = 0.0) -> StepCoeffs:
    """Stage coefficients in seconds.

    The reference forms RKB[s]*TimeStep in exact-rational TimeFrac arithmetic
    and evaluates it as Whole + Numer/Denom (src/infra/TimeMgr.cpp:382-391,
    747-766); for dt an integer number of seconds that is what
    ``_frac_seconds`` reproduces."""
    from fractions import Fraction

    def fs(fr: Fraction) -> float:
        x = Fraction(fr) * Fraction(dt).limit_denominator(10**9)
        whole = x.numerator // x.denominator if x >= 0 else -((-x.numerator) // x.denominator)
        rem = x - whole
        if rem == 0:
            return float(whole)
        return float(whole) + float(rem.numerator) / float(rem.denominator)

    c = StepCoeffs()
    c.Dt = fs(Fraction(1))
    if stepper == RK4:
        a = [Fraction(0), Fraction(1, 2), Fraction(1, 2), Fraction(1)]
        b = [Fraction(1, 6), Fraction(1, 3), Fraction(1, 3), Fraction(1, 6)]
        c.NStages = 4
        for s in range(4):
            c.A[s], c.B[s], c.StageTime[s] = fs(a[s]), fs(b[s]), t0 + fs(a[s])
    elif stepper == RK2:
        c.NStages = 2
        c.A[1], c.B[1] = fs(Fraction(1, 2)), fs(Fraction(1))
        c.StageTime[0], c.StageTime[1] = t0, t0 + fs(Fraction(1, 2))
    else:
        c.NStages = 1
        c.StageTime[0], c.StageTime[1] = t0, t0 + fs(Fraction(1))
    return c


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# OMEGA_B200_LIB: another build of the same library (profiling experiments); never a CPU fallback
LIB_PATH = os.environ.get("OMEGA_B200_LIB") or os.path.join(ROOT, "omega_b200", "csrc", "libomega_b200.so")


# the single-precision build of the same sources (Real = float; -DOMEGA_SINGLE_PRECISION)
LIB_PATH_SP = os.path.join(ROOT, "omega_b200", "csrc", "libomega_b200_sp.so")


def load_library(path: str = LIB_PATH) -> C.CDLL:
    """Load the CUDA library; raises (never falls back) when it is missing."""
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(the B200 path has no CPU fallback)")
    lib = C.CDLL(path)
    lib.omega_last_error.restype = C.c_char_p
    return lib
