"""The oracle against the REFERENCE ITSELF: /root/reference/src/tracer/MOM_tracer_advect.F90, unmodified, translated
statement by statement to C++ by oracle/f2c/f2cpp.py (no Fortran compiler exists in the image) and compiled with
g++ -O2 -ffp-contract=off into oracle/_ref/libref_f2c*.so.

* every committed golden vector (tests/golden/*.npz: all three schemes, both sweep orders, the five flux diagnostics,
  the offline optional arguments, non-symmetric memory, open boundaries on four sides) must be reproduced bit for bit
  by the translated reference -- so the goldens are the reference's answers, not only the oracle's;
* seeded synthetic cases beyond the goldens: hand-written oracle == translated reference, bitwise;
* the two things a translation cannot take from the source are run BOTH ways: what an unassigned automatic LOGICAL
  holds (the reference reads do_i one cell outside the range it sets, SURVEY hazard 7) and which argument MIN/MAX
  return on a +0/-0 tie (gfortran < 9: the first; >= 9: minsd/maxsd semantics) -- no compared output depends on either;
* a bounds-checked build runs every array reference of the translated reference inside its declared Fortran bounds.

The libraries are built here (where /root/reference exists) and travel to the GPU box; without either the tests skip."""
import importlib.util
import os

import numpy as np
import pytest

from mom6_b200 import synth
from mom6_b200.tracer_advect import tracer_advect_init
from oracle import f2c_ref, oracle

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
mg = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(mg)

pytestmark = pytest.mark.skipif(not f2c_ref.available(), reason="neither oracle/_ref/libref_f2c.so nor /root/reference present")

eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))
VARIANTS = ["", "lt", "mm", "chk"]


def test_translation_is_current_and_complete():
    """Where the reference source is present the translation is regenerated from it; it must contain the three
    subroutines with their reference line numbers and no hand edits (the file is generated, git-ignored)."""
    assert f2c_ref.build()
    inc = os.path.join(os.path.dirname(f2c_ref.__file__), "_ref", "advect_f2c.inc")
    if not os.path.exists(inc):
        pytest.skip("prebuilt libraries only (GPU box)")
    txt = open(inc).read()
    for name, line in (("advect_tracer", 52), ("advect_x", 329), ("advect_y", 705)):
        assert "MOM_tracer_advect.F90:%d  subroutine %s" % (line, name) in txt
    assert txt.count("// :") > 400      # one reference line number per translated statement


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("name", sorted(mg.CASES))
def test_translated_reference_reproduces_golden(name, variant):
    spec = mg.CASES[name]
    z = dict(np.load(os.path.join(HERE, "golden", name + ".npz")))
    c, Reg = mg.build_case(spec)
    for f in mg.GRID_FIELDS:
        getattr(c.G, f)[...] = z["G_" + f]
    for m, tr in enumerate(Reg.Tr):
        tr.t[...] = z["t_in_%02d" % m]
    for key, a in mg.obc_reservoirs(c.OBC):
        a[...] = z[key]
    ni, nj, nk, ntr, scheme, cfl, x_first, ncalls, diags, offline, kw = spec
    G, nk_ = c.G, c.GV.ke
    CS = tracer_advect_init(None, G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    uo, vo = np.zeros(G.shape_u(nk_)), np.zeros(G.shape_v(nk_))
    vol = z["vol_prev"].copy() if offline else None
    for call in range(ncalls):
        f2c_ref.advect_tracer(z["h_end"], z["uhtr"], z["vhtr"], c.OBC, c.dt, G, c.GV, None, CS, Reg, x_first_in=x_first,
                              vol_prev=vol, max_iter_in=2 if offline else None,
                              update_vol_prev=True if offline else None, uhr_out=uo, vhr_out=vo, variant=variant)
    for m, tr in enumerate(Reg.Tr):
        assert eq(G.interior(tr.t), G.interior(z["t_out_%02d" % m])), "tracer %d" % m
        if diags:
            for d in mg.DIAGS:
                cut = G.interior_u if d in ("ad_x", "ad2d_x") else G.interior_v if d in ("ad_y", "ad2d_y") else G.interior
                assert eq(cut(getattr(tr, d)), cut(z["%s_%02d" % (d, m)])), "%s of tracer %d" % (d, m)
    assert eq(G.interior_u(uo), G.interior_u(z["uhr_out"]))
    assert eq(G.interior_v(vo), G.interior_v(z["vhr_out"]))
    if offline:
        assert eq(G.interior(vol), G.interior(z["vol_prev_out"]))


def _pair(ni, nj, nk, ntr, scheme, variant="", bug=False, diags=False, ncalls=1, x_first=None, max_iter=None, **kw):
    c = synth.make_case(ni, nj, nk, ntr, **kw)
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme, "USE_HUYNH_STENCIL_BUG": bug}
    CS = tracer_advect_init(None, c.G, None, params, None, None)
    Ro, Rr = synth.bind_registry(c, diags=diags), synth.bind_registry(c, diags=diags)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    G, nk_ = c.G, c.GV.ke
    outs = [[np.zeros(G.shape_u(nk_)), np.zeros(G.shape_v(nk_))] for _ in range(2)]
    for _ in range(ncalls):
        oracle.advect_tracer(h, u, v, None, c.dt, G, c.GV, None, CS, Ro, x_first_in=x_first, max_iter_in=max_iter,
                             uhr_out=outs[0][0], vhr_out=outs[0][1])
        f2c_ref.advect_tracer(h, u, v, None, c.dt, G, c.GV, None, CS, Rr, x_first_in=x_first, max_iter_in=max_iter,
                              uhr_out=outs[1][0], vhr_out=outs[1][1], variant=variant)
    for m, (a, b) in enumerate(zip(Ro.Tr, Rr.Tr)):
        # the WHOLE arrays, halos included: the restatement follows the reference in what it leaves in the halos too
        assert eq(a.t, b.t), "tracer %d" % m
        if diags:
            for d in mg.DIAGS:
                assert eq(getattr(a, d), getattr(b, d)), "%s of tracer %d" % (d, m)
    assert eq(outs[0][0], outs[1][0]) and eq(outs[0][1], outs[1][1])


@pytest.mark.parametrize("scheme,bug", [("PLM", False), ("PPM", False), ("PPM:H3", False), ("PPM:H3", True)])
@pytest.mark.parametrize("x_first", [True, False])
@pytest.mark.parametrize("halo", [2, 3, 4, 6])
def test_oracle_equals_translated_reference_whole_arrays(scheme, bug, x_first, halo):
    if halo < (3 if scheme != "PLM" and not bug else 2):
        pytest.skip("halo narrower than the stencil")
    _pair(31, 23, 3, 4, scheme, bug=bug, x_first=x_first, halo=halo, cfl=0.9, diags=True, ncalls=2)


@pytest.mark.parametrize("variant", ["lt", "mm"])
@pytest.mark.parametrize("kw", [dict(cyclic_x=False), dict(symmetric=False), dict(cyclic_y=True), dict(land=False)])
def test_oracle_equals_translated_reference_topologies(variant, kw):
    _pair(26, 19, 2, 3, "PPM", variant=variant, cfl=1.4, diags=(variant != "lt"), **kw)


def test_oracle_equals_translated_reference_iteration_limits():
    for mi in (1, 2, 5):
        _pair(20, 16, 2, 3, "PLM", cfl=2.5, max_iter=mi)
        _pair(20, 16, 2, 3, "PPM", cfl=2.5, max_iter=mi)


def test_unassigned_logical_only_touches_outermost_diagnostic_faces():
    """With do_i's unassigned entries .true. instead of .false. (what a Fortran stack could hold) tracers, residuals and
    every diagnostic value away from the outermost faces of a sweep are unchanged."""
    c = synth.make_case(24, 18, 2, 3, cfl=1.2)
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": "PPM"}
    CS = tracer_advect_init(None, c.G, None, params, None, None)
    R0, R1 = synth.bind_registry(c, diags=True), synth.bind_registry(c, diags=True)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    f2c_ref.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CS, R0, variant="")
    f2c_ref.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CS, R1, variant="lt")
    G = c.G
    for a, b in zip(R0.Tr, R1.Tr):
        assert eq(a.t, b.t) and eq(a.advection_xy, b.advection_xy)
        assert eq(G.interior_u(a.ad_x), G.interior_u(b.ad_x)) and eq(G.interior_v(a.ad_y), G.interior_v(b.ad_y))


# ---- the loops either side of advect_tracer ---------------------------------------------------
@pytest.mark.parametrize("symmetric", [True, False])
def test_side_loops_equal_translated_reference_lines(symmetric):
    """oracle.accumulate_transports / oracle.column_diag (what the CUDA entry points of step.cu are checked against)
    against MOM_dynamics_split_RK2.F90:1045-1053 and MOM_tracer_registry.F90:744-767, 785-788, 795-798, 804-807
    translated statement by statement (bounds-checked build), whole arrays, bitwise."""
    from mom6_b200 import synth
    from oracle import oracle
    c = synth.make_case(23, 17, 6, 2, cfl=0.5, symmetric=symmetric)
    G, GV, nz = c.G, c.GV, c.GV.ke
    rng = np.random.default_rng(5)
    uh, vh = rng.standard_normal(G.shape_u(nz)), rng.standard_normal(G.shape_v(nz))
    u0, v0 = rng.standard_normal(G.shape_u(nz)), rng.standard_normal(G.shape_v(nz))
    ua, va, ub, vb = u0.copy(), v0.copy(), u0.copy(), v0.copy()
    oracle.accumulate_transports(ua, va, uh, vh, 37.5, G)
    f2c_ref.accumulate_transports(ub, vb, uh, vh, 37.5, G, GV)
    assert np.array_equal(ua.view(np.uint64), ub.view(np.uint64)) and np.array_equal(va.view(np.uint64), vb.view(np.uint64))
    assert not np.array_equal(ua, u0)
    t = rng.standard_normal(G.shape_h(nz))
    axy = rng.standard_normal(G.shape_h(nz))
    h = 40.0 * rng.random(G.shape_h(nz))
    h[:, G.pj(G.jsc), G.pi(G.isc)] = 0.0            # a vanished column and
    h[0, G.pj(G.jsc + 1), G.pi(G.isc + 1)] = 500.0  # one whose first layer already crosses 100 m
    for H_to_m in (0.75, 1.0):
        for kind, f3 in ((0, axy), (1, t), (2, t)):
            want, khi = f2c_ref.column_diag(kind, G, GV, f3, h, H_to_m)
            got = oracle.column_diag(kind, G, nz, t=t, h_diag=h, H_to_m=H_to_m, advection_xy=axy)
            assert np.array_equal(got.view(np.uint64), want.view(np.uint64)), kind
            if kind == 2:
                assert 1 <= khi <= nz
    # all layers above 100 m: khi = nz and zint_100m == zint
    hs = 0.1 * np.ones(G.shape_h(nz))
    w100, khi = f2c_ref.column_diag(2, G, GV, t, hs, 1.0)
    wz, _ = f2c_ref.column_diag(1, G, GV, t, hs, 1.0)
    assert khi == nz and np.array_equal(w100.view(np.uint64), wz.view(np.uint64))
    assert np.array_equal(oracle.column_diag(2, G, nz, t=t, h_diag=hs, H_to_m=1.0).view(np.uint64), w100.view(np.uint64))
