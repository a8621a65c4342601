"""tracer_hordiff, the along-layer diffusion iteration:
/root/reference/src/tracer/MOM_tracer_hor_diff.F90:539-607.

CPU: the hand-written oracle (oracle/advect_ref.cpp: ora_tracer_hordiff) against those lines of the UNMODIFIED
reference file translated statement by statement (oracle/f2c --range, bounds-checked), whole arrays, bitwise; plus the
properties diffusion must have.  GPU (marked): the CUDA entry point mom6adv_hordiff against the oracle, bitwise."""
import numpy as np
import pytest

from mom6_b200 import synth
from oracle import f2c_ref, oracle

eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))


def _case(ni=37, nj=29, nk=4, ntr=4, strength=0.3, **kw):
    c = synth.make_case(ni, nj, nk, ntr, cfl=0.4, **kw)
    kx, ky = synth.make_khdt(c, strength)
    nit = oracle.hordiff_num_itts(kx, ky, c.G)
    return c, kx, ky, nit


@pytest.mark.skipif(not f2c_ref.available(), reason="neither oracle/_ref nor /root/reference present")
@pytest.mark.parametrize("kw", [dict(), dict(cyclic_x=False), dict(symmetric=False), dict(cyclic_y=True), dict(tripolar_n=True)])
@pytest.mark.parametrize("ml", [None, (2, 3, 0.5), (2, 3, 0.0)])
def test_oracle_equals_translated_reference_lines(kw, ml):
    c, kx, ky, nit = _case(**kw)
    assert nit >= 2
    Ro, Rr = synth.bind_registry(c), synth.bind_registry(c)
    h = c.h_end.numpy()
    args = dict(Diffuse_ML_interior=ml is not None, nkml=ml[0] if ml else 0, nk_rho_varies=ml[1] if ml else 0,
                ML_KhTr_scale=ml[2] if ml else 0.0)
    oracle.tracer_hordiff(h, kx, ky, nit, c.G, c.GV, Ro, **args)
    f2c_ref.tracer_hordiff(h, kx, ky, nit, c.dt, c.G, c.GV, Rr, **args)
    for m in range(c.ntr):
        assert eq(Ro.Tr[m].t, Rr.Tr[m].t), "tracer %d" % m      # whole arrays, halos included


def test_diffusion_properties():
    """Constant tracers stay constant (bitwise); extrema do not grow; sum(T*h*areaT) is conserved to roundoff."""
    c, kx, ky, nit = _case(ni=30, nj=24, nk=3, ntr=3, cyclic_x=False)     # closed: the random khdt is not periodic
    G, h = c.G, c.h_end.numpy()
    Reg = synth.bind_registry(c)
    Reg.Tr[2].t[...] = 3.25
    before = [tr.t.copy() for tr in Reg.Tr]
    oracle.tracer_hordiff(h, kx, ky, nit, G, c.GV, Reg)
    assert eq(G.interior(Reg.Tr[2].t), G.interior(before[2]))
    ocean = G.interior(np.asarray(G.mask2dT)) > 0.5
    vol = G.interior(h) * G.interior(np.asarray(G.areaT))
    for m in (0, 1):
        a, b = G.interior(before[m]), G.interior(Reg.Tr[m].t)
        assert not eq(a, b)
        assert b[:, ocean].max() <= a[:, ocean].max() + 1e-12 and b[:, ocean].min() >= a[:, ocean].min() - 1e-12
        s0, s1 = oracle.efp_sum(a * vol)[1], oracle.efp_sum(b * vol)[1]
        assert abs(s1 - s0) <= 1e-11 * abs(s0)
