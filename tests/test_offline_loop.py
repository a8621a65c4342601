"""The iteration loop of offline_advection_ale:
/root/reference/src/tracer/MOM_offline_main.F90:254-312, 320-411, 414-415 with remaining_transport_sum (:615-650).

CPU: the hand-written oracle (oracle/advect_ref.cpp: ora_offline_advection_loop) against those lines of the UNMODIFIED
reference file translated statement by statement, running around the translated advect_tracer (oracle/f2c, bounds-checked):
tracers, h_pre, uhtr, vhtr, the convergence flag, the iteration count and every remaining_transport_sum, bitwise.
GPU (marked): mom6adv_offline_advect through the C ABI against the oracle, bitwise."""
import numpy as np
import pytest

from mom6_b200 import synth
from mom6_b200.tracer_advect import tracer_advect_init
from oracle import f2c_ref, oracle

eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))
UNITS = dict(H_to_kg_m2=1035.0, L_to_m=1.0, kg_m2_to_H=1.0 / 1035.0, m_to_L=1.0)


def _case(scheme="PPM", cfl=2.4, ni=37, nj=29, nk=3, ntr=3, **kw):
    c = synth.make_case(ni, nj, nk, ntr, cfl=cfl, **kw)
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    return c, CS


def _run(fn, c, CS, num_off_iter, x_before_y=True, min_residual=0.0, **kw):
    Reg = synth.bind_registry(c)
    h, u, v = c.h_end.numpy().copy(), c.uhtr.numpy().copy(), c.vhtr.numpy().copy()
    conv, nit, res = fn(h, u, v, c.dt, num_off_iter, c.G, c.GV, CS, Reg, x_before_y=x_before_y, min_residual=min_residual,
                        **UNITS, **kw)
    return Reg, h, u, v, conv, nit, res


@pytest.mark.skipif(not f2c_ref.available(), reason="neither oracle/_ref nor /root/reference present")
@pytest.mark.parametrize("kw", [dict(), dict(cyclic_x=False), dict(symmetric=False), dict(cyclic_y=True), dict(tripolar_n=True)])
@pytest.mark.parametrize("scheme,x_before_y", [("PPM", True), ("PLM", False), ("PPM:H3", True)])
def test_oracle_equals_translated_reference_loop(kw, scheme, x_before_y):
    c, CS = _case(scheme, **kw)
    a = _run(oracle.offline_advection_loop, c, CS, 8, x_before_y)
    b = _run(f2c_ref.offline_advection_loop, c, CS, 8, x_before_y)
    assert a[5] == b[5] and a[4] == b[4], (a[4:6], b[4:6])
    assert a[5] >= 2                                  # more than one trip through the loop
    assert eq(a[6], b[6]), (a[6], b[6])               # every remaining_transport_sum
    for m in range(c.ntr):
        assert eq(a[0].Tr[m].t, b[0].Tr[m].t), "tracer %d" % m
    assert eq(a[1], b[1]) and eq(a[2], b[2]) and eq(a[3], b[3])      # h_pre, uhtr, vhtr: whole arrays


@pytest.mark.skipif(not f2c_ref.available(), reason="neither oracle/_ref nor /root/reference present")
@pytest.mark.parametrize("scheme,x_before_y", [("PPM", True), ("PLM", False)])
def test_oracle_equals_translated_reference_loop_with_open_boundaries(scheme, x_before_y):
    """CS%OBC associated (:334): the regional basin with segments on all four sides, several trips."""
    c, OBC = synth.make_obc_case(cfl=1.8)
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    a = _run(oracle.offline_advection_loop, c, CS, 6, x_before_y, OBC=OBC)
    b = _run(f2c_ref.offline_advection_loop, c, CS, 6, x_before_y, OBC=OBC)
    n = _run(oracle.offline_advection_loop, c, CS, 6, x_before_y)
    assert a[5] == b[5] >= 2 and a[4] == b[4]
    assert eq(a[6], b[6])
    for m in range(c.ntr):
        assert eq(a[0].Tr[m].t, b[0].Tr[m].t), "tracer %d" % m
    assert eq(a[1], b[1]) and eq(a[2], b[2]) and eq(a[3], b[3])
    assert any(not eq(a[0].Tr[m].t, n[0].Tr[m].t) for m in range(c.ntr))      # the OBC matters


@pytest.mark.skipif(not f2c_ref.available(), reason="neither oracle/_ref nor /root/reference present")
def test_exit_paths():
    """Converged (remaining transport exactly 0), iteration cap, and the min_residual exit -- same in both."""
    c, CS = _case("PPM", cfl=0.4)
    for fn in (oracle.offline_advection_loop, f2c_ref.offline_advection_loop):
        Reg, h, u, v, conv, nit, res = _run(fn, c, CS, 6)
        assert conv and res[-1] == 0.0 and nit <= 2, (conv, nit, res)
    c, CS = _case("PPM", cfl=2.4)
    out = []
    for fn in (oracle.offline_advection_loop, f2c_ref.offline_advection_loop):
        Reg, h, u, v, conv, nit, res = _run(fn, c, CS, 2)
        assert not conv and nit == 2 and res[-1] > 0.0
        out.append(res)
        Reg, h, u, v, conv, nit, res = _run(fn, c, CS, 8, min_residual=0.5 * res[1])
        assert not conv and res[-1] < 0.5 * out[-1][1]
        out.append((nit, res))
    assert eq(out[0], out[2]) and out[1][0] == out[3][0] and eq(out[1][1], out[3][1])


def test_remaining_transport_sum_is_order_invariant_and_exact():
    """reproducing_sum's point: the value does not depend on the order of the columns, and for integers it is exact."""
    c, CS = _case("PLM", ni=24, nj=20, nk=2, ntr=1)
    G = c.G
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    r0 = oracle.remaining_transport_sum(u, v, h, G, c.GV, **UNITS)
    assert r0 > 0.0
    Gt, ht, ut, vt = synth.transposed_fields(c) if hasattr(synth, "transposed_fields") else (None, None, None, None)
    if Gt is not None:
        assert oracle.remaining_transport_sum(ut, vt, ht, Gt, c.GV, **UNITS) == r0
    # integers: |u| = 3 on every face of the computational domain, v = 0, unit scales -> 0.5 * (3 + 3) per cell and level
    u2, v2 = np.zeros_like(u), np.zeros_like(v)
    u2[:, G.pj(G.jsc):G.pj(G.jec) + 1, G.pI(G.isc) - 1:G.pI(G.iec) + 1] = 3.0
    n = (G.iec - G.isc + 1) * (G.jec - G.jsc + 1) * c.GV.ke
    assert oracle.remaining_transport_sum(u2, v2, h, G, c.GV) == 3.0 * n


# ---- the CUDA entry point (through the C ABI) against the oracle -------------------------------------------------
def _gpu_run(c, CS, num_off_iter, x_before_y=True, min_residual=0.0):
    from mom6_b200 import capi
    p = capi.Plan(c.G, c.GV, CS)
    p.set_ntr(c.ntr)
    R0 = synth.bind_registry(c)
    for m in range(c.ntr):
        p.upload_tracer(m, R0.Tr[m].t)
        p.bind_tracer(m, None, c.conc_underflow(m))
    p.upload_field(capi.F_H_END, c.h_end.numpy()); p.upload_field(capi.F_UHTR, c.uhtr.numpy()); p.upload_field(capi.F_VHTR, c.vhtr.numpy())
    conv, nit, res = p.offline_advect(c.dt, num_off_iter, x_before_y=x_before_y, min_residual=min_residual, **UNITS)
    nk = c.GV.ke
    T = []
    for m in range(c.ntr):
        o = np.zeros(c.G.shape_h(nk)); p.download_tracer(m, o); T.append(o)
    h, u, v = np.zeros(c.G.shape_h(nk)), np.zeros(c.G.shape_u(nk)), np.zeros(c.G.shape_v(nk))
    p.download_field(capi.F_H_END, h); p.download_field(capi.F_UHR, u); p.download_field(capi.F_VHR, v)
    p.destroy()
    return T, h, u, v, conv, nit, res


@pytest.mark.gpu
@pytest.mark.parametrize("kw", [dict(), dict(cyclic_x=False), dict(symmetric=False), dict(cyclic_y=True), dict(tripolar_n=True)])
@pytest.mark.parametrize("scheme,x_before_y", [("PPM", True), ("PLM", False), ("PPM:H3", True)])
def test_cuda_offline_loop_bitwise(kw, scheme, x_before_y):
    c, CS = _case(scheme, ni=70, nj=29, **kw)
    Ro, h, u, v, conv, nit, res = _run(oracle.offline_advection_loop, c, CS, 8, x_before_y)
    Tg, hg, ug, vg, convg, nitg, resg = _gpu_run(c, CS, 8, x_before_y)
    assert (conv, nit) == (convg, nitg) and nit >= 2
    assert eq(res, resg), (res, resg)                       # every remaining_transport_sum, bitwise
    G = c.G
    for m in range(c.ntr):
        assert eq(G.interior(Ro.Tr[m].t), G.interior(Tg[m])), "tracer %d" % m
    # h_pre, uhtr, vhtr: whole arrays less the outermost halo row either side (the plan's pass moves the rows
    # advect_tracer can read, the reference's pass_var / pass_vector all of them)
    assert eq(h[:, 1:-1, :], hg[:, 1:-1, :])
    assert eq(G.interior_u(u), G.interior_u(ug)) and eq(G.interior_v(v), G.interior_v(vg))
    assert eq(u[:, 1:-1, :], ug[:, 1:-1, :]) and eq(v[:, 2:-1, :], vg[:, 2:-1, :])


@pytest.mark.gpu
def test_cuda_offline_loop_exit_paths_and_refusals():
    from mom6_b200 import capi
    c, CS = _case("PPM", cfl=0.4)
    T, h, u, v, conv, nit, res = _gpu_run(c, CS, 6)
    Ro, ho, uo, vo, convo, nito, reso = _run(oracle.offline_advection_loop, c, CS, 6)
    assert conv and convo and nit == nito and eq(res, reso) and res[-1] == 0.0
    c, CS = _case("PPM", cfl=2.4)
    Ro, ho, uo, vo, convo, nito, reso = _run(oracle.offline_advection_loop, c, CS, 8, min_residual=1e30)
    T, h, u, v, conv, nit, res = _gpu_run(c, CS, 8, min_residual=1e30)
    assert (not conv) and nit == nito == 1 and eq(res, reso)
    for m in range(c.ntr):
        assert eq(c.G.interior(Ro.Tr[m].t), c.G.interior(T[m]))
    p = capi.Plan(c.G, c.GV, CS)
    p.set_ntr(1)
    with pytest.raises(capi.Mom6AdvError):
        p.offline_advect(c.dt, 4, off_ale_mod=2)          # the ALE step is not built: refused, not skipped
    with pytest.raises(capi.Mom6AdvError):
        p.offline_advect(c.dt, 0)
    p.destroy()


def _obc_offline_worker(scheme, x_before_y, q):
    """In a process of its own: this combination (offline loop + open boundaries on the device) has not run on a GPU."""
    try:
        from mom6_b200 import capi
        c, OBC = synth.make_obc_case(cfl=1.8)
        CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
        Ro, h, u, v, conv, nit, res = _run(oracle.offline_advection_loop, c, CS, 6, x_before_y, OBC=OBC)
        p = capi.Plan(c.G, c.GV, CS)
        p.set_ntr(c.ntr)
        R0 = synth.bind_registry(c)
        for m in range(c.ntr):
            p.upload_tracer(m, R0.Tr[m].t)
            p.bind_tracer(m, None, c.conc_underflow(m))
        p.set_obc(OBC)
        p.upload_field(capi.F_H_END, c.h_end.numpy()); p.upload_field(capi.F_UHTR, c.uhtr.numpy()); p.upload_field(capi.F_VHTR, c.vhtr.numpy())
        convg, nitg, resg = p.offline_advect(c.dt, 6, x_before_y=x_before_y, **UNITS)
        bad = []
        if (conv, nit) != (convg, nitg): bad.append("trips %s/%s" % (nitg, nit))
        if not eq(res, resg): bad.append("residuals")
        for m in range(c.ntr):
            o = np.zeros(c.G.shape_h(c.GV.ke))
            p.download_tracer(m, o)
            if not eq(c.G.interior(o), c.G.interior(Ro.Tr[m].t)): bad.append("tracer %d" % m)
        p.destroy()
        q.put("; ".join(bad))
    except Exception as e:
        q.put("raised %s: %s" % (type(e).__name__, str(e)[:300]))


@pytest.mark.gpu
@pytest.mark.xfail(strict=False, reason="offline loop with an associated OBC on the device: pieces verified separately, the combination never run on a GPU")
@pytest.mark.parametrize("scheme,x_before_y", [("PPM", True), ("PLM", False)])
def test_cuda_offline_loop_with_open_boundaries(scheme, x_before_y):
    import multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    pr = ctx.Process(target=_obc_offline_worker, args=(scheme, x_before_y, q))
    pr.start()
    try:
        msg = q.get(timeout=180)
    finally:
        pr.join(timeout=30)
        if pr.is_alive():
            pr.kill()
    assert msg == "", msg
