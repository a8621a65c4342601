"""tests/golden/row_*.npz -- answers of the two rows beside advect_tracer written by the reference's own lines
(tests/golden/make_golden_rows.py: MOM_tracer_hor_diff.F90:539-607 and MOM_offline_main.F90:270-415 + :634-648,
translated mechanically and run where /root/reference exists).  CPU: the oracle reproduces every file bit for bit.
GPU (marked): so does the CUDA path through the C ABI."""
import glob
import importlib.util
import os

import numpy as np
import pytest

from mom6_b200 import synth
from oracle import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_golden_rows", os.path.join(HERE, "golden", "make_golden_rows.py"))
mk = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(mk)

eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))


def _load(name):
    return dict(np.load(os.path.join(HERE, "golden", "row_%s.npz" % name)))


def test_every_listed_file_exists():
    have = {os.path.basename(p)[4:-4] for p in glob.glob(os.path.join(HERE, "golden", "row_*.npz"))}
    assert have == set(mk.HORDIFF) | set(mk.OFFLINE)


@pytest.mark.parametrize("name", sorted(mk.HORDIFF))
def test_oracle_reproduces_hordiff_golden(name):
    want = _load(name)
    got = mk.run_hordiff(oracle.tracer_hordiff, mk.HORDIFF[name], False)
    assert eq(got["inputs"], want["inputs"]), "the synthetic inputs are not the ones the file was written from"
    assert int(got["num_itts"][0]) == int(want["num_itts"][0]) >= 2
    for k in want:
        if k.startswith("t_"):
            assert eq(got[k], want[k]), k          # whole arrays, halos included


@pytest.mark.parametrize("name", sorted(mk.OFFLINE))
def test_oracle_reproduces_offline_golden(name):
    want = _load(name)
    got = mk.run_offline(oracle.offline_advection_loop, mk.OFFLINE[name])
    assert eq(got["inputs"], want["inputs"]), "the synthetic inputs are not the ones the file was written from"
    for k in ("converged", "trips"):
        assert int(got[k][0]) == int(want[k][0])
    assert int(want["trips"][0]) >= 2
    for k in want:
        if k.startswith("t_") or k in ("residuals", "h_pre", "uhtr", "vhtr"):
            assert eq(got[k], want[k]), k


# ---- the CUDA path against the same files --------------------------------------------------------------------
def _plan(c, CS):
    from mom6_b200 import capi
    p = capi.Plan(c.G, c.GV, CS)
    p.set_ntr(c.ntr)
    R0 = synth.bind_registry(c)
    for m in range(c.ntr):
        p.upload_tracer(m, R0.Tr[m].t)
        p.bind_tracer(m, None, c.conc_underflow(m))
    return p


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(mk.HORDIFF))
def test_cuda_reproduces_hordiff_golden(name):
    from mom6_b200.tracer_advect import tracer_advect_init
    want = _load(name)
    spec = mk.HORDIFF[name]
    c, kx, ky, nit = mk.hordiff_case(spec)
    ml = spec[5]
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": "PPM"}, None, None)
    p = _plan(c, CS)
    p.hordiff(kx, ky, nit, c.GV.H_subroundoff, h=c.h_end.numpy(), Diffuse_ML_interior=ml[0], nkml=ml[1], nk_rho_varies=ml[2],
              ML_KhTr_scale=ml[3])
    for m in range(c.ntr):
        o = np.zeros(c.G.shape_h(c.GV.ke))
        p.download_tracer(m, o)
        assert eq(c.G.interior(o), c.G.interior(want["t_%d" % m])), "tracer %d" % m
    p.destroy()


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(mk.OFFLINE))
def test_cuda_reproduces_offline_golden(name):
    from mom6_b200 import capi
    want = _load(name)
    spec = mk.OFFLINE[name]
    c, CS = mk.offline_case(spec)
    p = _plan(c, CS)
    p.upload_field(capi.F_H_END, c.h_end.numpy()); p.upload_field(capi.F_UHTR, c.uhtr.numpy()); p.upload_field(capi.F_VHTR, c.vhtr.numpy())
    conv, trips, res = p.offline_advect(c.dt, spec[7], x_before_y=spec[6], min_residual=spec[8], **mk.UNITS)
    assert int(conv) == int(want["converged"][0]) and trips == int(want["trips"][0])
    assert eq(res, want["residuals"])
    nk = c.GV.ke
    for m in range(c.ntr):
        o = np.zeros(c.G.shape_h(nk))
        p.download_tracer(m, o)
        assert eq(c.G.interior(o), c.G.interior(want["t_%d" % m])), "tracer %d" % m
    h, u, v = np.zeros(c.G.shape_h(nk)), np.zeros(c.G.shape_u(nk)), np.zeros(c.G.shape_v(nk))
    p.download_field(capi.F_H_END, h); p.download_field(capi.F_UHR, u); p.download_field(capi.F_VHR, v)
    assert eq(c.G.interior(h), c.G.interior(want["h_pre"]))
    assert eq(c.G.interior_u(u), c.G.interior_u(want["uhtr"])) and eq(c.G.interior_v(v), c.G.interior_v(want["vhtr"]))
    p.destroy()
