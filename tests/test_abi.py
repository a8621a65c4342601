"""The C-ABI library loads and exports every symbol include/mom6adv.h declares
(no compute calls: this runs without a GPU)."""
import ctypes
import os
import re

from mom6_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "mom6adv.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mom6adv_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_all_declared_symbols():
    lib_path = build.build()
    lib = ctypes.CDLL(lib_path)
    names = _declared_symbols()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), "symbol %s declared in mom6adv.h is not exported" % n
    assert sorted(capi.SYMBOLS) == names, "capi.SYMBOLS out of sync with mom6adv.h"


def test_version_and_no_cpu_fallback():
    l = capi.lib()
    assert b"sm_100a" in l.mom6adv_version()
    import torch
    if not torch.cuda.is_available():
        # without a device plan creation must fail loudly (status 4), never fall back
        from mom6_b200 import synth
        from mom6_b200.tracer_advect import tracer_advect_init
        import pytest
        c = synth.make_case(8, 8, 1, 1)
        CS = tracer_advect_init(None, c.G, None, {"DT": 900.0}, None, None)
        with pytest.raises(capi.Mom6AdvError) as e:
            capi.Plan(c.G, c.GV, CS)
        assert e.value.status == capi.ERR_CUDA


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "mom6_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in txt.replace("the oracle", "").replace("to the oracle", "") or f == "synth.py", \
                    "%s mentions the oracle" % f
