"""omega_b200/config.py: Omega's YAML configuration -> OmegaTendConfig, stepper, time step, halo width.  The values of
configs/Default.yml are what capi.TendConfig.default() hard-codes; where the reference tree is at hand the file itself
is read and compared (a CPU check; nothing on the GPU box reads /root/reference)."""
import os

import pytest

from omega_b200 import capi, config

DEFAULT_YML = "/root/reference/components/omega/configs/Default.yml"


def fields(c):
    return {n: getattr(c, n) for n, _ in capi.TendConfig._fields_}


def test_a_configuration_of_our_own():
    tree = {"Omega": {
        "TimeIntegration": {"TimeStepper": "RungeKutta4", "TimeStep": "0000_00:01:00"},
        "Decomp": {"HaloWidth": 4},
        "Advection": {"FluxThicknessType": "Upwind", "FluxTracerType": "Center"},
        "WindStress": {"InterpType": "Anisotropic"},
        "Tendencies": {"VelHyperDiffTendencyEnable": False, "ViscDel2": 250.0, "WindForcingTendencyEnable": True, "EddyDiff4": 3.5e9,
                       "UseCustomTendency": True, "ManufacturedSolutionTendency": True},
        "ManufacturedSolution": {"WavelengthX": 1.0e6, "Amplitude": 2.0},
        "Eos": {"EosType": "linear", "Linear": {"DRhoDT": -0.25}}}}
    got = config.from_dict(tree)
    want = capi.TendConfig.default(VelHyperDiffTendencyEnable=0, ViscDel2=250.0, WindForcingTendencyEnable=1, EddyDiff4=3.5e9,
                                   FluxThicknessType=1, WindStressInterpType=1, ManufacturedSolutionTendency=1,
                                   ManufacturedWavelengthX=1.0e6, ManufacturedAmplitude=2.0)
    assert fields(got["tend"]) == fields(want)
    assert got["stepper"] == capi.RK4 and got["dt"] == 60.0 and got["halo_width"] == 4 and got["ntimelevels"] == 2
    assert got["eos"] == dict(teos10=0, DRhodT=-0.25, DRhodS=0.8, RhoT0S0=1000.0)
    # one switch alone does not turn the manufactured terms on (Tendencies.cpp:48-62)
    tree["Omega"]["Tendencies"]["UseCustomTendency"] = False
    assert config.from_dict(tree)["tend"].ManufacturedSolutionTendency == 0


def test_time_intervals_and_unknown_choices():
    assert config.parse_time_interval("0000_00:10:00") == 600.0
    assert config.parse_time_interval("0001_01:00:30.5") == 86400 + 3600 + 30.5
    assert config.parse_time_interval("0000_00:00:0.2") == 0.2
    with pytest.raises(ValueError):
        config.parse_time_interval("ten minutes")
    with pytest.raises(ValueError):
        config.from_dict({"TimeIntegration": {"TimeStepper": "LeapFrog"}})
    with pytest.raises(ValueError):
        config.from_dict({"Advection": {"FluxThicknessType": "Centre"}})


def test_an_empty_configuration_is_the_default(tmp_path):
    p = tmp_path / "omega.yml"
    p.write_text("Omega:\n  Tendencies: {}\n")
    got = config.load(str(p))
    assert fields(got["tend"]) == fields(capi.TendConfig.default())
    assert got["stepper"] == capi.FB and got["dt"] == 600.0 and got["halo_width"] == 3


@pytest.mark.skipif(not os.path.exists(DEFAULT_YML), reason="the reference tree is not on this machine")
def test_default_yml_of_the_reference_equals_the_hard_coded_defaults():
    got = config.load(DEFAULT_YML)
    assert fields(got["tend"]) == fields(capi.TendConfig.default())
    assert got["stepper"] == capi.FB and got["dt"] == 600.0 and got["halo_width"] == 3 and got["ntimelevels"] == 2
    assert got["eos"] == dict(teos10=1, DRhodT=-0.2, DRhodS=0.8, RhoT0S0=1000.0)
