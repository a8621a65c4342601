"""GPU parity: CUDA advect_tracer (through the C ABI) vs the CPU oracle, bitwise."""
import numpy as np
import pytest

from mom6_b200 import synth
from mom6_b200.tracer_advect import tracer_advect_init, tracer_advect_end, advect_tracer
from oracle import oracle

pytestmark = pytest.mark.gpu


def _run_pair(ni, nj, nk, ntr, scheme, cfl=0.3, x_first=None, diags=False, ncalls=1, symmetric=True,
              cyclic_x=True, halo=4, bug=False, fast=2, sparse=1, sparse_cap=0, **kw):
    c = synth.make_case(ni, nj, nk, ntr, cfl=cfl, symmetric=symmetric, cyclic_x=cyclic_x, halo=halo, **kw)
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme, "USE_HUYNH_STENCIL_BUG": bug}
    CSo = tracer_advect_init(None, c.G, None, params, None, None)
    CSg = tracer_advect_init(None, c.G, None, params, None, None)
    CSg.gpu_options["fast"] = fast
    CSg.gpu_options["sparse"] = sparse
    CSg.gpu_options["sparse_cap"] = sparse_cap
    Ro = synth.bind_registry(c, diags=diags)
    Rg = synth.bind_registry(c, diags=diags)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    nk_ = c.GV.ke
    uo_o, vo_o = np.zeros(c.G.shape_u(nk_)), np.zeros(c.G.shape_v(nk_))
    uo_g, vo_g = np.zeros(c.G.shape_u(nk_)), np.zeros(c.G.shape_v(nk_))
    for n in range(ncalls):
        xf = x_first if x_first is not None else None
        nit_o = oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CSo, Ro, x_first_in=xf, uhr_out=uo_o, vhr_out=vo_o)
        nit_g = advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CSg, Rg, x_first_in=xf, uhr_out=uo_g, vhr_out=vo_g)
        assert nit_o == nit_g, (nit_o, nit_g)
    tracer_advect_end(CSg)
    return c, Ro, Rg, (uo_o, vo_o, uo_g, vo_g), nit_o


def _assert_bitwise(c, Ro, Rg, outs, diags=False):
    G = c.G
    for m, (a, b) in enumerate(zip(Ro.Tr, Rg.Tr)):
        ia, ib = G.interior(a.t), G.interior(b.t)
        assert np.array_equal(ia.view(np.uint64), ib.view(np.uint64)), \
            "tracer %d differs: max abs %g" % (m, np.abs(ia - ib).max())
        if diags:
            for n in ("ad_x", "ad_y", "ad2d_x", "ad2d_y", "advection_xy"):
                da, db = getattr(a, n), getattr(b, n)
                if n in ("ad_x", "ad2d_x"):
                    da, db = G.interior_u(da), G.interior_u(db)
                elif n in ("ad_y", "ad2d_y"):
                    da, db = G.interior_v(da), G.interior_v(db)
                else:
                    da, db = G.interior(da), G.interior(db)
                assert np.array_equal(da.view(np.uint64), db.view(np.uint64)), \
                    "diag %s of tracer %d differs: max abs %g" % (n, m, np.abs(da - db).max())
    uo_o, vo_o, uo_g, vo_g = outs
    assert np.array_equal(G.interior_u(uo_o).view(np.uint64), G.interior_u(uo_g).view(np.uint64))
    assert np.array_equal(G.interior_v(vo_o).view(np.uint64), G.interior_v(vo_g).view(np.uint64))


@pytest.mark.parametrize("scheme", ["PLM", "PPM", "PPM:H3"])
@pytest.mark.parametrize("x_first", [True, False])
@pytest.mark.parametrize("cfl", [0.3, 0.9])
def test_double_gyre_shape_bitwise(scheme, x_first, cfl):
    c, Ro, Rg, outs, nit = _run_pair(44, 40, 2, 10, scheme, cfl=cfl, x_first=x_first)
    _assert_bitwise(c, Ro, Rg, outs)


@pytest.mark.parametrize("scheme", ["PLM", "PPM", "PPM:H3"])
def test_diagnostics_bitwise(scheme):
    c, Ro, Rg, outs, nit = _run_pair(44, 40, 3, 4, scheme, cfl=0.9, diags=True)
    _assert_bitwise(c, Ro, Rg, outs, diags=True)


def test_tc1_shape_plm():
    c, Ro, Rg, outs, nit = _run_pair(10, 8, 8, 3, "PLM", cfl=0.5)
    _assert_bitwise(c, Ro, Rg, outs)


def test_repeated_calls():
    c, Ro, Rg, outs, nit = _run_pair(44, 40, 2, 10, "PPM", cfl=0.6, ncalls=5)
    _assert_bitwise(c, Ro, Rg, outs)


@pytest.mark.parametrize("symmetric", [True, False])
@pytest.mark.parametrize("cyclic_x", [True, False])
def test_memory_modes_and_topology(symmetric, cyclic_x):
    c, Ro, Rg, outs, nit = _run_pair(37, 29, 3, 3, "PPM", cfl=0.8, symmetric=symmetric, cyclic_x=cyclic_x)
    _assert_bitwise(c, Ro, Rg, outs)


def test_huynh_stencil_bug_and_halo3():
    c, Ro, Rg, outs, nit = _run_pair(40, 32, 2, 3, "PPM:H3", cfl=0.8, bug=True)
    _assert_bitwise(c, Ro, Rg, outs)
    c, Ro, Rg, outs, nit = _run_pair(40, 32, 2, 3, "PPM", cfl=0.8, halo=3)
    _assert_bitwise(c, Ro, Rg, outs)


def test_om4_1_slice_bitwise():
    # OM4_1 horizontal shape, a few levels (the oracle finishes in seconds)
    c, Ro, Rg, outs, nit = _run_pair(360, 210, 4, 6, "PPM", cfl=0.5)
    _assert_bitwise(c, Ro, Rg, outs)


# ---- streaming (fast) kernels vs generic kernels vs oracle -------------------------------------
@pytest.mark.parametrize("scheme", ["PLM", "PPM", "PPM:H3"])
@pytest.mark.parametrize("fast", [0, 1, 2])
def test_generic_and_streaming_kernels_both_bitwise(scheme, fast):
    c, Ro, Rg, outs, nit = _run_pair(44, 40, 2, 10, scheme, cfl=0.9, fast=fast)
    _assert_bitwise(c, Ro, Rg, outs)


@pytest.mark.parametrize("scheme", ["PLM", "PPM", "PPM:H3"])
@pytest.mark.parametrize("x_first", [True, False])
def test_many_x_tiles(scheme, x_first):
    # 300 columns = 3 warp tiles of the x kernel; 70 tracers = 2 tracer-table chunks
    c, Ro, Rg, outs, nit = _run_pair(300, 37, 2, 70 if scheme == "PPM" else 5, scheme, cfl=0.9, x_first=x_first)
    _assert_bitwise(c, Ro, Rg, outs)


@pytest.mark.parametrize("halo,scheme", [(2, "PLM"), (3, "PPM"), (5, "PPM"), (6, "PLM"), (7, "PPM:H3")])
def test_halo_widths(halo, scheme):
    c, Ro, Rg, outs, nit = _run_pair(50, 33, 2, 3, scheme, cfl=0.9, halo=halo)
    _assert_bitwise(c, Ro, Rg, outs)


def test_closed_domain_no_land():
    c, Ro, Rg, outs, nit = _run_pair(130, 31, 3, 4, "PPM", cfl=1.4, cyclic_x=False, land=False)
    assert nit >= 2
    _assert_bitwise(c, Ro, Rg, outs)


def test_resident_mode_matches_mirror_mode():
    """RESIDENT-mode calls (fields stay in HBM) give the same bits as MIRROR-mode calls."""
    from mom6_b200 import capi
    c = synth.make_case(60, 44, 3, 5, cfl=0.9)
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": "PPM"}
    CSo = tracer_advect_init(None, c.G, None, params, None, None)
    Ro = synth.bind_registry(c)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    for _ in range(3):
        oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CSo, Ro)
    CS = tracer_advect_init(None, c.G, None, params, None, None)
    p = capi.Plan(c.G, c.GV, CS)
    p.set_ntr(c.ntr)
    R0 = synth.bind_registry(c)
    for m in range(c.ntr):
        p.upload_tracer(m, R0.Tr[m].t)
        p.bind_tracer(m, None, c.conc_underflow(m))
    p.upload_field(capi.F_H_END, h); p.upload_field(capi.F_UHTR, u); p.upload_field(capi.F_VHTR, v)
    for _ in range(3):
        p.advect_resident(c.dt)
    for m in range(c.ntr):
        out = np.zeros_like(R0.Tr[m].t)
        p.download_tracer(m, out)
        assert np.array_equal(c.G.interior(out).view(np.uint64), c.G.interior(Ro.Tr[m].t).view(np.uint64))
    p.destroy()


@pytest.mark.parametrize("scheme", ["PLM", "PPM", "PPM:H3"])
@pytest.mark.parametrize("sparse", [0, 1])
@pytest.mark.parametrize("x_first", [True, False])
def test_residual_iterations_dense_and_sparse(scheme, sparse, x_first):
    # cfl 1.6: the limiter cuts many faces, three or more trips; the list-driven update must equal the dense one
    c, Ro, Rg, outs, nit = _run_pair(90, 41, 3, 6, scheme, cfl=1.6, x_first=x_first, sparse=sparse)
    assert nit >= 3
    _assert_bitwise(c, Ro, Rg, outs)


@pytest.mark.parametrize("max_iter", [1, 2, None])
@pytest.mark.parametrize("ntr", [3, 9])
def test_offline_arguments_bitwise(max_iter, ntr):
    """vol_prev / max_iter_in / update_vol_prev / uhr_out (offline tracer mode, :84-86,165-171,316-320).
    ntr = 9 is three pipeline chunks of the MIRROR call: vol_prev must be written back exactly once."""
    c = synth.make_case(70, 45, 3, ntr, cfl=1.5)
    G, nk = c.G, c.GV.ke
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": "PPM"}
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    rng = np.random.default_rng(5)
    vol0 = np.zeros(G.shape_h(nk))
    G.interior(vol0)[...] = G.interior(G.areaT) * G.interior(h) * rng.uniform(0.8, 1.2, G.interior(h).shape)
    res = []
    for impl in (oracle.advect_tracer, advect_tracer):
        CS = tracer_advect_init(None, G, None, params, None, None)
        R = synth.bind_registry(c)
        vol = vol0.copy()
        uo, vo = np.zeros(G.shape_u(nk)), np.zeros(G.shape_v(nk))
        uu, vv = u, v
        nits = []
        for call in range(2):      # second call continues from the residuals of the first, as the offline driver does
            nits.append(impl(h, uu, vv, None, c.dt, G, c.GV, None, CS, R, x_first_in=(call == 0), vol_prev=vol,
                             max_iter_in=max_iter, update_vol_prev=True, uhr_out=uo, vhr_out=vo))
            uu, vv = uo.copy(), vo.copy()
        if impl is advect_tracer:
            tracer_advect_end(CS)
        res.append((R, vol, uo, vo, nits))
    (Ro, vol_o, uo_o, vo_o, n_o), (Rg, vol_g, uo_g, vo_g, n_g) = res
    assert n_o == n_g
    _assert_bitwise(c, Ro, Rg, (uo_o, vo_o, uo_g, vo_g))
    assert np.array_equal(G.interior(vol_o).view(np.uint64), G.interior(vol_g).view(np.uint64))


@pytest.mark.parametrize("scheme", ["PLM", "PPM", "PPM:H3"])
@pytest.mark.parametrize("x_first", [True, False])
def test_hundred_consecutive_calls_bitwise(scheme, x_first):
    """double-gyre shape, 10 tracers, 100 consecutive calls, bitwise."""
    c, Ro, Rg, outs, nit = _run_pair(44, 40, 2, 10, scheme, cfl=0.45, x_first=x_first, ncalls=100)
    _assert_bitwise(c, Ro, Rg, outs)


@pytest.mark.parametrize("scheme", ["PLM", "PPM", "PPM:H3"])
@pytest.mark.parametrize("x_first", [True, False])
def test_open_boundaries_bitwise(scheme, x_first):
    """Associated OBC (:442-478,580-628,825-862,965-1014): specified and open segments on all four sides,
    tracer reservoirs and OBC_inflow_conc, inflow and outflow faces, an unassociated segment registry."""
    c, OBC = synth.make_obc_case()
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    nk = c.GV.ke
    res = []
    for impl in (oracle.advect_tracer, advect_tracer):
        CS = tracer_advect_init(None, c.G, None, params, None, None)
        R = synth.bind_registry(c, diags=True)
        uo, vo = np.zeros(c.G.shape_u(nk)), np.zeros(c.G.shape_v(nk))
        nits = [impl(h, u, v, OBC, c.dt, c.G, c.GV, None, CS, R, x_first_in=x_first, uhr_out=uo, vhr_out=vo) for _ in range(2)]
        if impl is advect_tracer:
            tracer_advect_end(CS)
        res.append((R, uo, vo, nits))
    (Ro, uo_o, vo_o, n_o), (Rg, uo_g, vo_g, n_g) = res
    assert n_o == n_g
    # the OBC must matter: the same call without it gives another answer
    CS = tracer_advect_init(None, c.G, None, params, None, None)
    Rn = synth.bind_registry(c)
    oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CS, Rn, x_first_in=x_first)
    assert any(not np.array_equal(c.G.interior(a.t), c.G.interior(b.t)) for a, b in zip(Rn.Tr, Ro.Tr))
    _assert_bitwise(c, Ro, Rg, (uo_o, vo_o, uo_g, vo_g), diags=True)


@pytest.mark.parametrize("scheme", ["PLM", "PPM"])
@pytest.mark.parametrize("x_first", [True, False])
@pytest.mark.parametrize("cap", [8, 300])
def test_residual_list_overflow_falls_back_to_dense(scheme, x_first, cap):
    """Residual trips skip the mass-field tiles without residual transport; when the active-cell list overflows
    (forced here by a tiny capacity) the skipped tiles are completed and the dense kernels take over."""
    c, Ro, Rg, outs, nit = _run_pair(300, 41, 3, 6, scheme, cfl=1.6, x_first=x_first, sparse_cap=cap)
    assert nit >= 3
    _assert_bitwise(c, Ro, Rg, outs)


def test_sparse_trips_wide_grid_many_tiles():
    # 700 columns = 23 warp tiles of the mass kernels per row: residual faces on tile edges, skipped tiles in between
    c, Ro, Rg, outs, nit = _run_pair(700, 37, 2, 4, "PPM", cfl=0.9)
    assert nit >= 2
    _assert_bitwise(c, Ro, Rg, outs)
    c, Ro, Rg, outs, nit = _run_pair(700, 37, 2, 4, "PPM", cfl=0.9, x_first=False, cyclic_x=False)
    _assert_bitwise(c, Ro, Rg, outs)


# ---- either side of advect_tracer: transport accumulation and column diagnostics on the device ----
@pytest.mark.parametrize("symmetric", [True, False])
def test_transport_accumulation_and_column_diagnostics_bitwise(symmetric):
    """mom6adv_accumulate_transports / zero_transports / column_diag against the numpy restatement of
    MOM_dynamics_split_RK2.F90:1045-1052, MOM.F90:1465-1466 and MOM_tracer_registry.F90:725-815; then an advect
    call fed by the device-accumulated transports against the oracle fed by the host-accumulated ones."""
    from mom6_b200 import capi
    c = synth.make_case(60, 44, 6, 4, cfl=0.6, symmetric=symmetric)
    G, nz = c.G, c.GV.ke
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": "PPM"}
    CS = tracer_advect_init(None, G, None, params, None, None)
    p = capi.Plan(G, c.GV, CS)
    p.set_ntr(c.ntr)
    Rg = synth.bind_registry(c)
    adv_xy = [np.zeros(G.shape_h(nz)) for _ in range(c.ntr)]
    for m in range(c.ntr):
        p.upload_tracer(m, Rg.Tr[m].t)
        p.bind_tracer(m, None, c.conc_underflow(m), advection_xy=adv_xy[m] if m == 1 else None)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    # three dynamics steps of transports that add up to (roughly) the case's uhtr / vhtr
    rng = np.random.default_rng(7)
    dts = [300.0, 450.0, 150.0]
    uhs = [u / 900.0 * (1.0 + 0.1 * rng.standard_normal(u.shape)) for _ in dts]
    vhs = [v / 900.0 * (1.0 + 0.1 * rng.standard_normal(v.shape)) for _ in dts]
    uacc, vacc = np.zeros_like(u), np.zeros_like(v)
    p.upload_field(capi.F_H_END, h)
    p.upload_field(capi.F_UHTR, u); p.upload_field(capi.F_VHTR, v)      # something non-zero to be zeroed
    p.zero_transports()
    for uh, vh, dt in zip(uhs, vhs, dts):
        oracle.accumulate_transports(uacc, vacc, uh, vh, dt, G)
        p.accumulate_transports(uh, vh, dt)
    ug, vg = np.full_like(u, 9.0), np.full_like(v, 9.0)
    p.download_field(capi.F_UHTR, ug); p.download_field(capi.F_VHTR, vg)
    assert np.array_equal(ug.view(np.uint64), uacc.view(np.uint64))
    assert np.array_equal(vg.view(np.uint64), vacc.view(np.uint64))
    # advect with the accumulated transports on both sides
    CSo = tracer_advect_init(None, G, None, params, None, None)
    Ro = synth.bind_registry(c, diags=False)
    Ro.Tr[1].advection_xy = np.zeros(G.shape_h(nz))
    oracle.advect_tracer(h, uacc, vacc, None, c.dt, G, c.GV, None, CSo, Ro)
    p.advect_resident(c.dt)
    for m in range(c.ntr):
        out = np.zeros_like(Rg.Tr[m].t)
        p.download_tracer(m, out)
        assert np.array_equal(G.interior(out).view(np.uint64), G.interior(Ro.Tr[m].t).view(np.uint64)), m
    # column diagnostics of the advected state
    h_diag = h * (1.0 + 0.05 * rng.standard_normal(h.shape))
    H_to_m = 0.5
    for m in range(c.ntr):
        for kind in (capi.CD_ZINT, capi.CD_ZINT_100M):
            ref = oracle.column_diag(kind, G, nz, t=Ro.Tr[m].t, h_diag=h_diag, H_to_m=H_to_m)
            out = np.full(G.shape_h(), 7.0)
            p.column_diag(m, kind, h_diag, H_to_m, out)
            assert np.array_equal(out.view(np.uint64), ref.view(np.uint64)), (m, kind, np.abs(out - ref).max())
    ref = oracle.column_diag(capi.CD_ADV_XY_2D, G, nz, advection_xy=Ro.Tr[1].advection_xy)
    out = np.full(G.shape_h(), 7.0)
    p.column_diag(1, capi.CD_ADV_XY_2D, None, 1.0, out)
    assert np.array_equal(out.view(np.uint64), ref.view(np.uint64))
    assert np.abs(ref).max() > 0.0
    p.destroy()


@pytest.mark.parametrize("scheme", ["PLM", "PPM", "PPM:H3"])
def test_cuda_equals_translated_reference(scheme):
    """CUDA path against the UNMODIFIED reference file translated mechanically to C++ (oracle/f2c, libraries prebuilt in
    oracle/_ref/ and shipped with the snapshot): tracers, residual transports and all five diagnostics, bitwise."""
    from oracle import f2c_ref
    if not f2c_ref.available():
        pytest.skip("oracle/_ref/libref_f2c.so not present")
    c = synth.make_case(44, 40, 3, 6, cfl=0.9)
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}
    CSr = tracer_advect_init(None, c.G, None, params, None, None)
    CSg = tracer_advect_init(None, c.G, None, params, None, None)
    Rr, Rg = synth.bind_registry(c, diags=True), synth.bind_registry(c, diags=True)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    nk_ = c.GV.ke
    uo_r, vo_r = np.zeros(c.G.shape_u(nk_)), np.zeros(c.G.shape_v(nk_))
    uo_g, vo_g = np.zeros(c.G.shape_u(nk_)), np.zeros(c.G.shape_v(nk_))
    for _ in range(2):
        f2c_ref.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CSr, Rr, uhr_out=uo_r, vhr_out=vo_r)
        advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CSg, Rg, uhr_out=uo_g, vhr_out=vo_g)
    tracer_advect_end(CSg)
    _assert_bitwise(c, Rr, Rg, (uo_r, vo_r, uo_g, vo_g), diags=True)


# ---- topologies and shapes the first round never ran end to end on the device ------------------------------------------
@pytest.mark.parametrize("kw", [dict(tripolar_n=True), dict(cyclic_y=True), dict(cyclic_x=False, cyclic_y=True)])
@pytest.mark.parametrize("scheme", ["PPM", "PLM"])
def test_tripolar_and_cyclic_y_bitwise(kw, scheme):
    """1-PE group pass through the tripolar north fold and through REENTRANT_Y, with transport crossing the seam."""
    c, Ro, Rg, outs, nit = _run_pair(48, 36, 3, 5, scheme, cfl=0.9, ncalls=2, **kw)
    _assert_bitwise(c, Ro, Rg, outs)


@pytest.mark.parametrize("scheme", ["PPM", "PLM"])
def test_transposed_case_bitwise(scheme):
    """The i <-> j transposed problem (y sweep first) on the device equals the transposed 1-PE answer of the original."""
    c = synth.make_case(70, 45, 3, 5, cfl=0.9)
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}
    CSo = tracer_advect_init(None, c.G, None, params, None, None)
    Ro = synth.bind_registry(c)
    G2, h2, u2, v2, R2 = synth.transposed_case(c, synth.bind_registry(c))
    oracle.advect_tracer(c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy(), None, c.dt, c.G, c.GV, None, CSo, Ro, x_first_in=True)
    CSg = tracer_advect_init(None, G2, None, params, None, None)
    advect_tracer(h2, u2, v2, None, c.dt, G2, c.GV, None, CSg, R2, x_first_in=False)
    tracer_advect_end(CSg)
    for m in range(Ro.ntr):
        a = np.ascontiguousarray(np.swapaxes(c.G.interior(Ro.Tr[m].t), -1, -2))
        b = np.ascontiguousarray(G2.interior(R2.Tr[m].t))
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), "tracer %d" % m


def test_om4_1_full_shape_bitwise():
    """BASELINE.json configs[2] at its full shape: 360 x 210 x 75, T/S + 20 passive tracers, PPM."""
    ni, nj, nk, ntr, scheme = synth.CONFIGS["OM4_1"]
    c, Ro, Rg, outs, nit = _run_pair(ni, nj, nk, ntr, scheme, cfl=0.45)
    assert nit >= 2
    _assert_bitwise(c, Ro, Rg, outs)


# ---- flux diagnostics on the streaming path (DIAG instantiations, list-driven trips, k-ordered ad2d) ---------------------
def _partial_diags(c, which):
    """Registry in which tracer m carries the diagnostics named in which[m] (a tuple of attribute names)."""
    Reg = synth.bind_registry(c)
    G, nk = c.G, c.GV.ke
    shapes = {"ad_x": G.shape_u(nk), "ad_y": G.shape_v(nk), "ad2d_x": G.shape_u(), "ad2d_y": G.shape_v(),
              "advection_xy": G.shape_h(nk)}
    for m, names in which.items():
        for n in names:
            setattr(Reg.Tr[m], n, np.full(shapes[n], np.nan))
    return Reg


@pytest.mark.parametrize("scheme", ["PPM", "PLM", "PPM:H3"])
@pytest.mark.parametrize("x_first", [True, False])
@pytest.mark.parametrize("cfl", [0.5, 1.6])
def test_partial_diagnostics_on_the_streaming_path(scheme, x_first, cfl):
    """T carries ad_x / ad_y (the routine T_adx, T_ady of production diag_tables, MOM_tracer_registry.F90:406-407), S
    advection_xy only, one passive tracer the 2-D sums, one everything, the others nothing: the registry is split into
    a DIAG launch and a plain launch of the same kernels; cfl 1.6 adds list-driven residual trips."""
    which = {0: ("ad_x", "ad_y"), 1: ("advection_xy",), 3: ("ad2d_x", "ad2d_y"),
             5: ("ad_x", "ad_y", "ad2d_x", "ad2d_y", "advection_xy"), 6: ("ad_y",)}
    c = synth.make_case(150, 61, 3, 8, cfl=cfl)
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}
    CSo = tracer_advect_init(None, c.G, None, params, None, None)
    CSg = tracer_advect_init(None, c.G, None, params, None, None)
    Ro, Rg = _partial_diags(c, which), _partial_diags(c, which)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    for _ in range(2):
        no = oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CSo, Ro, x_first_in=x_first)
        ng = advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CSg, Rg, x_first_in=x_first)
        assert no == ng
    if cfl > 1.0:
        assert ng >= 3, "the case was meant to exercise the list-driven residual trips"
    tracer_advect_end(CSg)
    G = c.G
    eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))
    for m in range(c.ntr):
        assert eq(G.interior(Ro.Tr[m].t), G.interior(Rg.Tr[m].t)), "tracer %d" % m
        for n in which.get(m, ()):
            cut = G.interior_u if n in ("ad_x", "ad2d_x") else G.interior_v if n in ("ad_y", "ad2d_y") else G.interior
            a, b = cut(getattr(Ro.Tr[m], n)), cut(getattr(Rg.Tr[m], n))
            assert eq(a, b), "%s of tracer %d: max abs %g" % (n, m, np.abs(a - b).max())
            assert np.abs(a).max() > 0.0
