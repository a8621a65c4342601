"""The exact-by-construction rewrites of the streaming kernels (fast_common.cuh), compiled as HOST code
and compared bit for bit with the literal restatement of MOM_tracer_advect.F90:427-431,517-577
(internal.cuh) on millions of stencils: +-0, equal neighbours, denormals, masked cells, both flow
directions, PLM / PPM / PPM:H3, and the exact /3 and /6 sequences against true division."""
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ARITH = os.path.join(ROOT, "tests", "arith")


def _gxx():
    for c in ("/usr/bin/g++", shutil.which("g++")):
        if c and os.path.exists(c):
            return c
    pytest.skip("no g++")


def test_streaming_rewrites_bitwise(tmp_path):
    exe = str(tmp_path / "check_fast_arith")
    cmd = [_gxx(), "-O2", "-ffp-contract=off", "-I" + os.path.join(ARITH, "stub"),
           "-I" + os.path.join(ROOT, "mom6_b200", "csrc"), "-o", exe, os.path.join(ARITH, "check_fast_arith.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-3000:]
    r = subprocess.run([exe, "1500000"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:]
    assert " 0 mismatches" in r.stdout
