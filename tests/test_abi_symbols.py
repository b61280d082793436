"""CPU checks of the drop-in boundary: the C-ABI library loads without a GPU and
exports every function include/omega_b200.h declares; struct layouts used by
the ctypes mirror match the header's field order."""
import ctypes
import os
import re

import pytest

from omega_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "omega_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(omega_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    lib = ctypes.CDLL(capi.LIB_PATH)
    names = declared_functions()
    assert len(names) >= 35
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.omega_b200_abi_version() == 1


def test_ctypes_struct_fields_follow_the_header():
    src = open(HEADER).read()

    def fields(struct):
        body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (struct, struct), src, re.S).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        out = []
        for decl in body.split(";"):
            decl = decl.strip()
            if not decl:
                continue
            decl = re.sub(r"^(const\s+)?(int|double|OmegaReal)\s*", "", decl)
            for part in decl.split(","):
                out.append(re.sub(r"[\*\s]|\[\d+\]", "", part))
        return out

    for struct, cls in (("OmegaMeshDesc", capi.MeshDesc), ("OmegaTendConfig", capi.TendConfig),
                        ("OmegaAuxArrays", capi.AuxArrays), ("OmegaTendArrays", capi.TendArrays),
                        ("OmegaStateArrays", capi.StateArrays), ("OmegaStepCoeffs", capi.StepCoeffs)):
        assert fields(struct) == [n for n, _ in cls._fields_], struct


def test_missing_library_fails_loudly(tmp_path):
    with pytest.raises(RuntimeError):
        capi.load_library(str(tmp_path / "nope.so"))
