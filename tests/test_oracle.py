"""CPU tests of the oracle itself.

The reference has no golden vectors for advect_tracer, so the
oracle is held to the reference's own *relational* standards: layout
invariance (test.layout), memory-mode invariance (test.grid), threading
invariance (test.openmp), dimensional rescaling (test.dim.*), plus the
scheme's analytic properties (constancy, monotonicity, conservation).
"""
import numpy as np
import pytest

from mom6_b200 import synth, decomp
from mom6_b200.types import TracerRegistry, register_tracer
from mom6_b200.tracer_advect import tracer_advect_init
from oracle import oracle

SCHEMES = ["PLM", "PPM", "PPM:H3"]


def _cs(c, scheme, bug=False):
    return tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme,
                                                "USE_HUYNH_STENCIL_BUG": bug}, None, None)


def _run(c, scheme, Reg=None, **kw):
    CS = _cs(c, scheme)
    Reg = Reg if Reg is not None else synth.bind_registry(c)
    nit = oracle.advect_tracer(c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy(), None, c.dt, c.G, c.GV, None,
                               CS, Reg, **kw)
    return Reg, nit


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


@pytest.mark.parametrize("scheme", SCHEMES)
def test_constant_tracer_is_preserved(scheme):
    c = synth.make_case(30, 24, 3, 1, cfl=0.7)
    Reg = TracerRegistry()
    register_tracer(np.full(c.G.shape_h(3), 7.25), Reg)
    _run(c, scheme, Reg)
    wet = c.G.interior(c.G.mask2dT) > 0
    t = c.G.interior(Reg.Tr[0].t)[:, wet]
    assert np.abs(t - 7.25).max() < 1e-12


@pytest.mark.parametrize("scheme", SCHEMES)
@pytest.mark.parametrize("cfl", [0.3, 0.9])
def test_monotone_and_finite(scheme, cfl):
    c = synth.make_case(44, 40, 2, 10, cfl=cfl)
    Reg0 = synth.bind_registry(c)
    Reg, nit = _run(c, scheme)
    wet = c.G.interior(c.G.mask2dT) > 0
    for a, b in zip(Reg0.Tr, Reg.Tr):
        t0, t1 = c.G.interior(a.t)[:, wet], c.G.interior(b.t)[:, wet]
        assert np.isfinite(t1).all()
        span = max(t0.max() - t0.min(), 1.0)
        assert t1.max() <= t0.max() + 1e-12 * span
        assert t1.min() >= t0.min() - 1e-12 * span


@pytest.mark.parametrize("scheme", SCHEMES)
def test_conservation(scheme):
    """sum(T*hprev) over the (cyclic-x, closed-y) domain is conserved to roundoff."""
    c = synth.make_case(40, 32, 3, 6, cfl=0.8)
    G, nk = c.G, c.GV.ke
    # start volume == what advect_tracer reconstructs (:165-171); passed through vol_prev
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    vol = np.zeros(G.shape_h(nk))
    ui, vi = G.interior_u(u), G.interior_v(v)
    hp = G.interior(G.areaT) * G.interior(h) + ((ui[..., 1:] - ui[..., :-1]) + (vi[:, 1:, :] - vi[:, :-1, :]))
    hp = np.maximum(0.0, hp)
    G.interior(vol)[...] = hp
    Reg0 = synth.bind_registry(c)
    Reg = synth.bind_registry(c)
    CS = _cs(c, scheme)
    oracle.advect_tracer(h, u, v, None, c.dt, G, c.GV, None, CS, Reg, vol_prev=vol, update_vol_prev=True)
    for a, b in zip(Reg0.Tr, Reg.Tr):
        s0 = oracle.efp_sum(G.interior(a.t) * hp)[1]
        s1 = oracle.efp_sum(G.interior(b.t) * G.interior(vol))[1]
        scale = oracle.efp_sum(np.abs(G.interior(a.t)) * hp)[1]
        assert abs(s1 - s0) <= 1e-11 * scale


def test_zero_transport_is_identity():
    c = synth.make_case(20, 16, 2, 3)
    Reg0 = synth.bind_registry(c)
    Reg = synth.bind_registry(c)
    CS = _cs(c, "PPM")
    z_u, z_v = np.zeros_like(c.uhtr.numpy()), np.zeros_like(c.vhtr.numpy())
    nit = oracle.advect_tracer(c.h_end.numpy(), z_u, z_v, None, c.dt, c.G, c.GV, None, CS, Reg)
    assert nit == 1
    for a, b in zip(Reg0.Tr, Reg.Tr):
        assert np.array_equal(_bits(c.G.interior(a.t)), _bits(c.G.interior(b.t)))


def test_ntr_zero_returns_immediately():
    c = synth.make_case(12, 10, 2, 0)
    assert oracle.advect_tracer(c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy(), None, c.dt, c.G, c.GV, None,
                                _cs(c, "PLM"), TracerRegistry()) == 0


def test_max_iter_default_tc1():
    # hazard 10: 2*ceil(3600/900)+1 = 9
    from mom6_b200.tracer_advect import max_iter_default
    c = synth.make_case(10, 8, 8, 3)
    assert max_iter_default(3600.0, _cs(c, "PLM")) == 9


@pytest.mark.parametrize("scheme", SCHEMES)
def test_threading_invariance(scheme):
    c = synth.make_case(36, 28, 6, 4, cfl=0.9)
    R1 = synth.bind_registry(c, diags=True)
    R4 = synth.bind_registry(c, diags=True)
    CS = _cs(c, scheme)
    args = (c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy(), None, c.dt, c.G, c.GV, None, CS)
    oracle.advect_tracer(*args, R1, nthreads=1)
    oracle.advect_tracer(*args, R4, nthreads=4)
    for a, b in zip(R1.Tr, R4.Tr):
        assert np.array_equal(_bits(a.t), _bits(b.t))
        for n in ("ad_x", "ad_y", "ad2d_x", "ad2d_y", "advection_xy"):
            assert np.array_equal(_bits(getattr(a, n)), _bits(getattr(b, n)))


@pytest.mark.parametrize("scheme", SCHEMES)
def test_symmetric_vs_nonsymmetric_memory(scheme):
    cs = synth.make_case(33, 27, 2, 4, cfl=0.8, symmetric=True)
    cn = synth.make_case(33, 27, 2, 4, cfl=0.8, symmetric=False)
    Rs, n1 = _run(cs, scheme)
    Rn, n2 = _run(cn, scheme)
    assert n1 == n2
    for a, b in zip(Rs.Tr, Rn.Tr):
        assert np.array_equal(_bits(cs.G.interior(a.t)), _bits(cn.G.interior(b.t)))


def _split(c, layout, diags=False):
    """Decompose a global case into per-rank oracle inputs."""
    Gg = c.G
    Regg = synth.bind_registry(c, diags=False)
    ranks = []
    for pj in range(layout[1]):
        for pi in range(layout[0]):
            G = decomp.local_grid(Gg, layout, (pi, pj))
            Reg = TracerRegistry()
            for tr in Regg.Tr:
                register_tracer(decomp.cut_h(Gg, G, tr.t), Reg, underflow_conc=tr.conc_underflow)
            ranks.append(dict(G=G, GV=c.GV, Reg=Reg,
                              h_end=decomp.cut_h(Gg, G, c.h_end.numpy()),
                              uhtr=decomp.cut_u(Gg, G, c.uhtr.numpy()),
                              vhtr=decomp.cut_v(Gg, G, c.vhtr.numpy())))
    return ranks


@pytest.mark.parametrize("scheme", SCHEMES)
@pytest.mark.parametrize("layout", [(2, 1), (1, 2), (2, 2), (3, 2)])
@pytest.mark.parametrize("x_first", [True, False])
def test_layout_invariance(scheme, layout, x_first):
    """N-PE answer == 1-PE answer bitwise (the reference's test.layout standard).

    y-first runs use a land-free grid: the reference's domore_v rescan (:221-225)
    only looks at a PE's own columns [is,ie] while the y-first sweep also covers
    the halo columns [is-stencil, ie+stencil] (:276-277), so a PE whose own part of
    a row is all land skips halo columns that the owning PE advects -- a genuine
    layout dependence of the reference in that corner (findings)."""
    c = synth.make_case(36, 30, 2, 4, cfl=0.9, land=x_first)
    R1, nit1 = _run(c, scheme, x_first_in=x_first)
    ranks = _split(c, layout)
    CS = _cs(c, scheme)
    nits = oracle.advect_tracer_multi(ranks, c.G.Domain, CS, c.dt, x_first_in=x_first)
    for m in range(c.ntr):
        glob = np.zeros(c.G.shape_h(c.GV.ke))
        for r in ranks:
            decomp.paste_interior(c.G, r["G"], glob, r["Reg"].Tr[m].t)
        assert np.array_equal(_bits(c.G.interior(glob)), _bits(c.G.interior(R1.Tr[m].t))), \
            "tracer %d differs between layout %s and 1 PE" % (m, layout)


@pytest.mark.parametrize("scheme", SCHEMES)
def test_power_of_two_rescaling(scheme):
    """test.dim.*: rescaling H by 2^a and L^2 by 2^(2b) leaves tracer bits unchanged."""
    c = synth.make_case(32, 26, 2, 4, cfl=0.8)
    R0, n0 = _run(c, scheme)
    a, b2 = 2.0 ** 11, 2.0 ** 22
    c2 = synth.make_case(32, 26, 2, 4, cfl=0.8)
    c2.G.areaT = c2.G.areaT * b2
    c2.G.IareaT = c2.G.IareaT / b2
    c2.GV.Angstrom_H *= a
    c2.GV.H_subroundoff *= a
    c2.h_end = c2.h_end * a
    c2.uhtr = c2.uhtr * (a * b2)
    c2.vhtr = c2.vhtr * (a * b2)
    R2, n2 = _run(c2, scheme)
    assert n0 == n2
    for x, y in zip(R0.Tr, R2.Tr):
        assert np.array_equal(_bits(c.G.interior(x.t)), _bits(c.G.interior(y.t)))


def test_offline_arguments_roundtrip():
    """max_iter_in=1 + uhr_out/vhr_out: a second call on the residuals finishes the job."""
    c = synth.make_case(30, 24, 2, 3, cfl=1.5)
    G, nk = c.G, c.GV.ke
    CS = _cs(c, "PPM")
    Rfull, nfull = _run(c, "PPM")
    assert nfull >= 2
    uo, vo = np.zeros(G.shape_u(nk)), np.zeros(G.shape_v(nk))
    R = synth.bind_registry(c)
    n1 = oracle.advect_tracer(c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy(), None, c.dt, G, c.GV, None, CS, R,
                              max_iter_in=1, uhr_out=uo, vhr_out=vo)
    assert n1 == 1
    assert np.abs(G.interior_u(uo)).max() > 0 or np.abs(G.interior_v(vo)).max() > 0


def test_metrics():
    a = np.array([1.0, -2.5, 3.0e10, 1e-20])
    limbs, val = oracle.efp_sum(a)
    assert abs(val - a.sum()) <= 1e-6
    assert oracle.efp_sum(a[::-1].copy())[0] == limbs      # order invariant
    assert oracle.bitcount_chksum(np.array([1.0, 3.0])) == bin(np.float64(1.0).view(np.uint64)).count("1") + \
        bin(np.float64(3.0).view(np.uint64)).count("1")


# ---- open boundaries (:442-478,580-628,825-862,965-1014) ------------------------------------------------------
@pytest.mark.parametrize("scheme", SCHEMES)
def test_obc_thread_count_invariance_and_effect(scheme):
    c, OBC = synth.make_obc_case()
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    out = []
    for nthreads, obc in ((1, OBC), (4, OBC), (1, None)):
        Reg = synth.bind_registry(c, diags=True)
        oracle.advect_tracer(h, u, v, obc, c.dt, c.G, c.GV, None, _cs(c, scheme), Reg, nthreads=nthreads)
        out.append(Reg)
    for a, b in zip(out[0].Tr, out[1].Tr):
        assert np.array_equal(_bits(a.t), _bits(b.t)) and np.array_equal(_bits(a.ad2d_x), _bits(b.ad2d_x))
        assert np.isfinite(c.G.interior(a.t)).all()
    assert any(not np.array_equal(c.G.interior(a.t), c.G.interior(b.t)) for a, b in zip(out[0].Tr, out[2].Tr))


def test_obc_uniform_tracer_with_matching_reservoir_stays_uniform():
    """T == const everywhere and every reservoir / inflow concentration == the same const: the prescribed inflow
    fluxes carry exactly that concentration, so the tracer stays uniform wherever cells keep some volume."""
    c, OBC = synth.make_obc_case(ntr=2)
    for sg in OBC.segment:
        if sg.tr_Reg is None:
            continue
        for t in sg.tr_Reg:
            t.ntr_index = min(t.ntr_index, 2)
            t.OBC_inflow_conc = 7.0
            if t.tres is not None:
                t.tres[...] = 7.0
    Reg = synth.bind_registry(c)
    for tr in Reg.Tr:
        tr.t[...] = 7.0
    oracle.advect_tracer(c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy(), OBC, c.dt, c.G, c.GV, None, _cs(c, "PPM"), Reg)
    wet = c.G.interior(c.G.mask2dT) > 0
    for tr in Reg.Tr:
        assert np.allclose(c.G.interior(tr.t)[:, wet], 7.0, rtol=1e-9)


def test_side_loops_numpy_restatement_matches_literal_loops():
    """oracle.accumulate_transports / column_diag (vectorised over i, j) against literal triple loops written from
    MOM_dynamics_split_RK2.F90:1045-1052 and MOM_tracer_registry.F90:745-811 on a tiny grid."""
    from mom6_b200 import synth
    c = synth.make_case(9, 7, 5, 2, cfl=0.5)
    G, nz = c.G, c.GV.ke
    rng = np.random.default_rng(3)
    uh, vh = rng.standard_normal(G.shape_u(nz)), rng.standard_normal(G.shape_v(nz))
    u0, v0 = rng.standard_normal(G.shape_u(nz)), rng.standard_normal(G.shape_v(nz))
    ua, va = u0.copy(), v0.copy()
    oracle.accumulate_transports(ua, va, uh, vh, 37.5, G)
    ub, vb = u0.copy(), v0.copy()
    Isq, Ieq, Jsq, Jeq = G.isc - 1, G.iec, G.jsc - 1, G.jec
    for k in range(nz):
        for j in range(G.jsc - 2, G.jec + 3):
            for I in range(Isq - 2, Ieq + 3):
                ub[k, G.pj(j), G.pI(I)] = ub[k, G.pj(j), G.pI(I)] + uh[k, G.pj(j), G.pI(I)] * 37.5
        for J in range(Jsq - 2, Jeq + 3):
            for i in range(G.isc - 2, G.iec + 3):
                vb[k, G.pJ(J), G.pi(i)] = vb[k, G.pJ(J), G.pi(i)] + vh[k, G.pJ(J), G.pi(i)] * 37.5
    assert np.array_equal(ua.view(np.uint64), ub.view(np.uint64)) and np.array_equal(va.view(np.uint64), vb.view(np.uint64))
    t = rng.standard_normal(G.shape_h(nz))
    h = 40.0 * rng.random(G.shape_h(nz))
    h[:, G.pj(G.jsc), G.pi(G.isc)] = 0.0            # a vanished column and
    h[0, G.pj(G.jsc + 1), G.pi(G.isc + 1)] = 500.0  # one whose first layer already crosses 100 m
    H_to_m = 0.75
    w1 = np.zeros(G.shape_h()); w2 = np.zeros(G.shape_h())
    frac = np.zeros(G.shape_h(nz)); khi = 0
    zbot = np.zeros(G.shape_h())
    for k in range(nz):
        anyp = False
        for j in range(G.jsc, G.jec + 1):
            for i in range(G.isc, G.iec + 1):
                pj, pi = G.pj(j), G.pi(i)
                ztop = zbot[pj, pi]
                zbot[pj, pi] = ztop + h[k, pj, pi] * H_to_m
                if zbot[pj, pi] <= 100.0:
                    frac[k, pj, pi] = 1.0
                elif ztop < 100.0:
                    frac[k, pj, pi] = (100.0 - ztop) / (zbot[pj, pi] - ztop)
                else:
                    frac[k, pj, pi] = 0.0
                anyp = anyp or frac[k, pj, pi] > 0
        if anyp:
            khi = k + 1
    for k in range(nz):
        for j in range(G.jsc, G.jec + 1):
            for i in range(G.isc, G.iec + 1):
                pj, pi = G.pj(j), G.pi(i)
                w1[pj, pi] = w1[pj, pi] + (h[k, pj, pi] * H_to_m) * t[k, pj, pi]
                if k < khi:
                    w2[pj, pi] = w2[pj, pi] + frac[k, pj, pi] * ((h[k, pj, pi] * H_to_m) * t[k, pj, pi])
    a1 = oracle.column_diag(1, G, nz, t=t, h_diag=h, H_to_m=H_to_m)
    a2 = oracle.column_diag(2, G, nz, t=t, h_diag=h, H_to_m=H_to_m)
    assert np.array_equal(a1.view(np.uint64), w1.view(np.uint64))
    assert np.array_equal(a2.view(np.uint64), w2.view(np.uint64))


# ---- index transposition (the reference's rotation standard, .testing/Makefile:384,455,596, in its simplest form) -----
@pytest.mark.parametrize("scheme", ["PLM", "PPM", "PPM:H3"])
@pytest.mark.parametrize("cyclic_x", [True, False])
def test_index_transposition_invariance(scheme, cyclic_x):
    """i <-> j, uhtr <-> vhtr, REENTRANT_X <-> REENTRANT_Y and the sweep order flipped give the transposed answer bit for
    bit: advect_x and advect_y are the same mathematics (their coded asymmetries -- SURVEY hazard 5 -- never change a
    value when volumes stay positive)."""
    c = synth.make_case(34, 27, 3, 5, cfl=0.9, cyclic_x=cyclic_x)
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}
    CS = tracer_advect_init(None, c.G, None, params, None, None)
    Reg = synth.bind_registry(c)
    G2, h2, u2, v2, Reg2 = synth.transposed_case(c, Reg)
    G, nk = c.G, c.GV.ke
    uo, vo = np.zeros(G.shape_u(nk)), np.zeros(G.shape_v(nk))
    uo2, vo2 = np.zeros(G2.shape_u(nk)), np.zeros(G2.shape_v(nk))
    for _ in range(2):
        n1 = oracle.advect_tracer(c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy(), None, c.dt, G, c.GV, None, CS, Reg,
                                  x_first_in=True, uhr_out=uo, vhr_out=vo)
        n2 = oracle.advect_tracer(h2, u2, v2, None, c.dt, G2, c.GV, None, CS, Reg2, x_first_in=False, uhr_out=uo2, vhr_out=vo2)
        assert n1 == n2
    eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))
    for m in range(Reg.ntr):
        assert eq(np.swapaxes(G.interior(Reg.Tr[m].t), -1, -2), G2.interior(Reg2.Tr[m].t)), "tracer %d" % m
    assert eq(np.swapaxes(G.interior_u(uo), -1, -2), G2.interior_v(vo2))
    assert eq(np.swapaxes(G.interior_v(vo), -1, -2), G2.interior_u(uo2))


@pytest.mark.parametrize("layout", [(1, 2), (2, 2), (3, 1)])
def test_tripolar_fold_layout_invariance(layout):
    """End to end through the tripolar north fold (src/framework/MOM_domains.F90:184-191): the emulated multi-PE run equals
    the 1-PE run bitwise, with transport crossing the fold row."""
    from mom6_b200 import decomp
    from mom6_b200.types import TracerRegistry, register_tracer
    c = synth.make_case(24, 18, 2, 3, cfl=0.9, tripolar_n=True)
    assert np.abs(c.vhtr.numpy()[:, c.G.pJ(c.G.jec), :]).max() > 0.0
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": "PPM"}
    CS = tracer_advect_init(None, c.G, None, params, None, None)
    R1 = synth.bind_registry(c)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CS, R1)
    R0 = synth.bind_registry(c)
    ranks = []
    for pj in range(layout[1]):
        for pi in range(layout[0]):
            G = decomp.local_grid(c.G, layout, (pi, pj))
            Reg = TracerRegistry()
            for tr in R0.Tr:
                register_tracer(decomp.cut_h(c.G, G, tr.t), Reg, underflow_conc=tr.conc_underflow)
            ranks.append({"G": G, "GV": c.GV, "Reg": Reg, "h_end": decomp.cut_h(c.G, G, h),
                          "uhtr": decomp.cut_u(c.G, G, u), "vhtr": decomp.cut_v(c.G, G, v)})
    import dataclasses
    dom = dataclasses.replace(c.G.Domain, layout=tuple(layout))
    oracle.advect_tracer_multi(ranks, dom, CS, c.dt)
    for r in ranks:
        G = r["G"]
        for m in range(c.ntr):
            want = decomp.cut_h(c.G, G, R1.Tr[m].t)
            assert np.array_equal(G.interior(want).view(np.uint64), G.interior(r["Reg"].Tr[m].t).view(np.uint64)), (layout, m)


@pytest.mark.parametrize("scheme", ["PPM", "PLM"])
def test_obc_basin_layouts(scheme):
    """The regional basin of the open-boundary tests, run per PE (every PE with its own clipped segments,
    decomp.cut_obc = setup_segment_indices) against the 1-PE run.  Equal bit for bit on the y-split layout, for y-first
    sweeps, and after one trip.  With x-first sweeps and a second trip an x-split layout differs from the 1-PE run at a
    few cells under the northern boundary -- and does so at the same cells with the OBC unassociated: not an
    open-boundary effect but a layout dependence of the per-PE algorithm (or of the group-pass emulation) in this basin,
    recorded here so that a change of it is noticed."""
    from mom6_b200.types import TracerRegistry, register_tracer

    def run(layout, use_obc, translated=False, **kw):
        c, OBC = synth.make_obc_case()
        if not use_obc:
            OBC = None
        CS = _cs(c, scheme)
        h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
        R1 = synth.bind_registry(c)
        oracle.advect_tracer(h, u, v, OBC, c.dt, c.G, c.GV, None, CS, R1, **kw)
        Rg = synth.bind_registry(c)
        ranks = []
        for pj in range(layout[1]):
            for pi in range(layout[0]):
                G = decomp.local_grid(c.G, layout, (pi, pj))
                Reg = TracerRegistry()
                for tr in Rg.Tr:
                    register_tracer(decomp.cut_h(c.G, G, tr.t), Reg, underflow_conc=tr.conc_underflow)
                ranks.append(dict(G=G, GV=c.GV, Reg=Reg, h_end=decomp.cut_h(c.G, G, h), uhtr=decomp.cut_u(c.G, G, u),
                                  vhtr=decomp.cut_v(c.G, G, v), OBC=decomp.cut_obc(OBC, c.G, G)))
        if translated:
            from oracle import f2c_ref
            f2c_ref.advect_tracer_multi(ranks, c.G.Domain, CS, c.dt, **kw)
        else:
            oracle.advect_tracer_multi(ranks, c.G.Domain, CS, c.dt, **kw)
        cells = set()
        for m in range(c.ntr):
            glob = np.zeros(c.G.shape_h(c.GV.ke))
            for r in ranks:
                decomp.paste_interior(c.G, r["G"], glob, r["Reg"].Tr[m].t)
            for (k, j, i) in np.argwhere(_bits(c.G.interior(glob)) != _bits(c.G.interior(R1.Tr[m].t))):
                cells.add((int(j), int(i)))
        return cells

    for layout in ((1, 3), (1, 4)):                      # y-split: equal whatever the order and the number of trips
        assert run(layout, True) == set() and run(layout, True, x_first_in=False) == set()
    for layout in ((2, 2), (3, 1)):                      # x-split: equal for y-first sweeps and after one trip
        assert run(layout, True, x_first_in=False) == set() and run(layout, True, max_iter_in=1) == set()
    for use_obc in (True, False):
        assert run((1, 2), use_obc) == set()
        assert run((1, 2), use_obc, x_first_in=False) == set()
        assert run((2, 1), use_obc, x_first_in=False) == set()
        assert run((2, 1), use_obc, max_iter_in=1) == set()
    # the same per-PE runs with the TRANSLATED REFERENCE as every PE's advect_tracer: identical cell sets, so the layout
    # dependence is in the reference's statements, not in the restatement
    from oracle import f2c_ref
    if f2c_ref.available():
        for use_obc in (True, False):
            assert run((2, 1), use_obc, translated=True) == run((2, 1), use_obc) != set()
            assert run((1, 2), use_obc, translated=True) == set()
            assert run((2, 1), use_obc, translated=True, x_first_in=False) == set()
    with_obc, without = run((2, 1), True), run((2, 1), False)
    # the same handful of cells under the northern boundary, OBC or not
    assert without and with_obc and len(with_obc | without) <= 12 and len(with_obc & without) >= 4
    assert all(31 <= j <= 32 for (j, i) in with_obc | without)
