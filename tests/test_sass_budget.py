"""Instruction budget of the streaming sweeps, read from the built library with cuobjdump (no GPU needed).

DESIGN.md section 4: on B200 an FP64 instruction holds its scheduler's issue port for two cycles, so a sweep costs
(non-FP64) + 2 x (FP64) issue cycles per cell and tracer.  This round's work was to shrink the non-FP64 part; the
numbers below are the state it reached (scripts/sass_loop.py --hot), with a little slack, so that a change which
silently brings back per-tracer predicate rebuilding, sign flips, spills or the per-lane copy loop fails here, on the
CPU, before anyone spends GPU time on it."""
import os
import re
import shutil
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "mom6_b200", "libmom6adv_b200.so")
TOOL = os.path.join(ROOT, "scripts", "sass_loop.py")


def _run(*args):
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    if not os.path.exists(LIB):
        pytest.skip("library not built")
    r = subprocess.run([sys.executable, TOOL, LIB] + list(args), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout


def _counts(out):
    m = re.search(r"instructions (\d+)\s+fp64 (\d+)\s+other (\d+)", out)
    ops = {k: int(v) for k, v in re.findall(r"^\s+([A-Z0-9_.]+)\s+(\d+)\s*$", out, re.M)}
    return int(m.group(1)), int(m.group(2)), int(m.group(3)), ops


def test_x_sweep_hot_loop_budget():
    # one iteration of the tracer loop = 4 cells of one tracer; rare regions (denormal /6, overshoot fix, underflow) skipped
    n, f64, other, ops = _counts(_run("k_x_fastILi1ELb0E", "--hot"))
    assert f64 <= 160, (n, f64, other)            # 155: the reference's arithmetic, each quantity once
    assert other <= 200, (n, f64, other)          # 186 (first generation: 318)
    assert ops.get("DSETP", 0) <= 10              # no per-tracer rebuilding of the tracer-independent flags
    assert "LDL" not in ops and "STL" not in ops  # no spills in the loop


@pytest.mark.parametrize("scheme", [0, 1, 2])
def test_sweeps_do_not_spill_and_producer_is_uniform(scheme):
    n, f64, other, ops = _counts(_run("k_y_ringILi%dE" % scheme, "--all"))
    assert ops.get("LDL", 0) + ops.get("STL", 0) <= 6, ops          # at most a couple of spilled words outside the march
    assert ops.get("UBLKCP", 0) >= 7                                  # bulk copies issued from uniform registers ...
    assert ops.get("ELECT", 0) <= 2 and ops.get("R2UR", 0) <= 12, ops   # ... not through a per-lane ELECT / R2UR loop
    nx, fx, ox, opsx = _counts(_run("k_x_fastILi%dELb0E" % scheme, "--all"))
    assert opsx.get("LDL", 0) + opsx.get("STL", 0) == 0, opsx


def test_hordiff_iteration_kernel_stages_through_shared_memory():
    """k_hd_tile (tracer_hordiff's iteration): the tracer tiles arrive by asynchronous
    global->shared copies (LDGSTS + LDGDEPBAR groups), the stencil is read from shared memory, nothing spills, and the
    register count keeps three 512-thread blocks on an SM (65536 / (3 * 512) = 42)."""
    n, f64, other, ops = _counts(_run("k_hd_tile", "--all"))
    assert ops.get("LDGSTS", 0) >= 2 and ops.get("LDGDEPBAR", 0) >= 2, ops
    assert ops.get("LDS.64", 0) + ops.get("LDS", 0) >= 5, ops
    assert "LDL" not in ops and "STL" not in ops, ops
    if not shutil.which("cuobjdump"):
        pytest.skip("cuobjdump not on PATH")
    r = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, timeout=300)
    m = re.search(r"Function _Z9k_hd_tile[^\n]*:\s*\n\s*REG:(\d+)", r.stdout)
    assert m, "k_hd_tile not found in the resource usage listing"
    assert int(m.group(1)) <= 42, m.group(1)
