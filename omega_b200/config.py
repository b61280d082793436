"""Omega's YAML configuration (configs/Default.yml: the `Omega:` tree) -> what the library takes.

The groups on the hot path and where the reference reads them:
  Tendencies:            src/ocn/Tendencies.cpp:123-213 (readTendConfig)      -> OmegaTendConfig flags and coefficients
  Advection:             src/ocn/AuxiliaryState.cpp:267-290                    -> FluxThicknessType / FluxTracerType
  WindStress:            src/ocn/AuxiliaryState.cpp:292-308                    -> WindStressInterpType
  ManufacturedSolution:  src/ocn/CustomTendencyTerms.cpp:83-102                -> wavelengths, amplitude
  TimeIntegration:       src/timeStepping/TimeStepper.cpp:226-305 (init1)      -> stepper type, time step in seconds
  Decomp:                src/base/Decomp.cpp:1704-1740                         -> HaloWidth
  State:                 src/ocn/OceanState.cpp:30-60                          -> NTimeLevels
  Eos:                   src/ocn/Eos.cpp:60-110                                -> EosType, linear coefficients
A key that is absent keeps the value of configs/Default.yml (capi.TendConfig.default), as a missing optional key does
upstream; an unknown choice raises, as upstream aborts."""
from __future__ import annotations

import re
from fractions import Fraction

from . import capi

_STEPPERS = {"Forward-Backward": capi.FB, "RungeKutta4": capi.RK4, "RungeKutta2": capi.RK2}
_TEND_BOOLS = ("ThicknessFluxTendencyEnable", "PVTendencyEnable", "KETendencyEnable", "SSHTendencyEnable", "VelDiffTendencyEnable",
               "VelHyperDiffTendencyEnable", "WindForcingTendencyEnable", "BottomDragTendencyEnable", "TracerHorzAdvTendencyEnable",
               "TracerDiffTendencyEnable", "TracerHyperDiffTendencyEnable")
_TEND_REALS = ("ViscDel2", "ViscDel4", "DivFactor", "Density0", "BottomDragCoeff", "EddyDiff2", "EddyDiff4")


def parse_time_interval(text) -> float:
    """`DDDD_HH:MM:SS(.fff)` (TimeInterval from a string, src/infra/TimeMgr.cpp) -> seconds, exactly when the
    seconds field has a few decimals"""
    m = re.fullmatch(r"\s*(\d+)_(\d+):(\d+):(\d+(?:\.\d*)?)\s*", str(text))
    if not m:
        raise ValueError(f"not a time interval: {text!r} (DDDD_HH:MM:SS)")
    d, h, mi = (int(m.group(i)) for i in (1, 2, 3))
    return float(Fraction(m.group(4)) + 60 * (mi + 60 * (h + 24 * d)))


def _choice(value, table, what):
    if value not in table:
        raise ValueError(f"unknown {what}: {value!r} (one of {sorted(table)})")
    return table[value]


def from_dict(tree: dict) -> dict:
    """tree: the parsed YAML, with or without the top-level `Omega:` key"""
    o = tree.get("Omega", tree)
    kw = {}
    t = o.get("Tendencies", {})
    for k in _TEND_BOOLS:
        if k in t:
            kw[k] = int(bool(t[k]))
    for k in _TEND_REALS:
        if k in t:
            kw[k] = float(t[k])
    # Tendencies.cpp:48-62: the manufactured terms need both switches
    kw["ManufacturedSolutionTendency"] = int(bool(t.get("UseCustomTendency", False)) and bool(t.get("ManufacturedSolutionTendency", False)))
    adv = o.get("Advection", {})
    flux = {"Center": 0, "Upwind": 1}
    if "FluxThicknessType" in adv:
        kw["FluxThicknessType"] = _choice(adv["FluxThicknessType"], flux, "FluxThicknessType")
    if "FluxTracerType" in adv:
        kw["FluxTracerType"] = _choice(adv["FluxTracerType"], flux, "FluxTracerType")
    if "InterpType" in o.get("WindStress", {}):
        kw["WindStressInterpType"] = _choice(o["WindStress"]["InterpType"], {"Isotropic": 0, "Anisotropic": 1}, "WindStress InterpType")
    ms = o.get("ManufacturedSolution", {})
    for k, name in (("WavelengthX", "ManufacturedWavelengthX"), ("WavelengthY", "ManufacturedWavelengthY"), ("Amplitude", "ManufacturedAmplitude")):
        if k in ms:
            kw[name] = float(ms[k])
    out = {"tend": capi.TendConfig.default(**kw)}
    ti = o.get("TimeIntegration", {})
    out["stepper"] = _choice(ti.get("TimeStepper", "Forward-Backward"), _STEPPERS, "TimeStepper")
    out["dt"] = parse_time_interval(ti.get("TimeStep", "0000_00:10:00"))
    out["halo_width"] = int(o.get("Decomp", {}).get("HaloWidth", 3))
    out["ntimelevels"] = int(o.get("State", {}).get("NTimeLevels", 2))
    eos = o.get("Eos", {})
    lin = eos.get("Linear", {})
    out["eos"] = dict(teos10=_choice(str(eos.get("EosType", "teos10")).lower(), {"teos10": 1, "teos-10": 1, "linear": 0}, "EosType"),
                      DRhodT=float(lin.get("DRhoDT", -0.2)), DRhodS=float(lin.get("DRhoDS", 0.8)), RhoT0S0=float(lin.get("RhoT0S0", 1000.0)))
    return out


def load(path: str) -> dict:
    import yaml
    with open(path) as f:
        return from_dict(yaml.safe_load(f))
