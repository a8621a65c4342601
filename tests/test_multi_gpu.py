"""N-GPU answer == 1-PE answer, bitwise (the reference's test.layout standard,
.testing/Makefile:383,454,595): one rank per GPU, NCCL halo exchange."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _worker(rank, world, port, layout, scheme, cfl, q, pull=1):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from mom6_b200 import synth, decomp, multi
        from mom6_b200.types import TracerRegistry, register_tracer
        from mom6_b200.tracer_advect import tracer_advect_init, tracer_advect_end, advect_tracer
        from oracle import oracle
        c = synth.make_case(96, 70, 3, 5, cfl=cfl)
        params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}
        # 1-PE oracle answer
        CSo = tracer_advect_init(None, c.G, None, params, None, None)
        Ro = synth.bind_registry(c)
        h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
        nit_o = 0
        for _ in range(2):
            nit_o = oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CSo, Ro)
        # this rank's block on its GPU
        pe = multi.rank_to_pe(rank, layout)
        G = decomp.local_grid(c.G, layout, pe)
        R0 = synth.bind_registry(c)
        Reg = TracerRegistry()
        for tr in R0.Tr:
            register_tracer(decomp.cut_h(c.G, G, tr.t), Reg, underflow_conc=tr.conc_underflow)
        CS = tracer_advect_init(None, G, None, params, None, None)
        CS.gpu_options["pull"] = pull      # 1: peer-memory loads over NVLink, 0: pack -> NCCL send/recv -> unpack
        hl, ul, vl = decomp.cut_h(c.G, G, h), decomp.cut_u(c.G, G, u), decomp.cut_v(c.G, G, v)
        nit = 0
        for _ in range(2):
            nit = advect_tracer(hl, ul, vl, None, c.dt, G, c.GV, None, CS, Reg)
        ok = (nit == nit_o)
        for m in range(c.ntr):
            want = decomp.cut_h(c.G, G, Ro.Tr[m].t)
            ok = ok and np.array_equal(G.interior(want).view(np.uint64), G.interior(Reg.Tr[m].t).view(np.uint64))
        tracer_advect_end(CS)
        q.put((rank, bool(ok), nit, nit_o))
    except Exception as e:      # report instead of leaving the parent to wait for its queue time-out
        q.put((rank, False, -1, "worker raised %s: %s" % (type(e).__name__, str(e)[:300])))
    finally:
        dist.destroy_process_group()


def _run(world, layout, scheme, cfl, pull=1):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [ctx.Process(target=_worker, args=(r, world, port, layout, scheme, cfl, q, pull)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert [r[:2] for r in res] == [(r, True) for r in range(world)], res


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("layout", [(1, 2), (2, 1)])
@pytest.mark.parametrize("scheme,cfl", [("PPM", 0.9), ("PLM", 0.5)])
def test_two_gpu_layout_invariance(layout, scheme, cfl):
    _run(2, layout, scheme, cfl)


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
def test_two_gpu_nccl_send_recv_path():
    _run(2, (1, 2), "PPM", 0.9, pull=0)


@pytest.mark.skipif(_ngpu() < 4, reason="needs 4 GPUs")
@pytest.mark.parametrize("layout", [(2, 2), (1, 4)])
def test_four_gpu_layout_invariance(layout):
    _run(4, layout, "PPM", 1.4)


@pytest.mark.skipif(_ngpu() < 8, reason="needs 8 GPUs")
@pytest.mark.parametrize("layout", [(2, 4), (1, 8)])
def test_eight_gpu_layout_invariance(layout):
    _run(8, layout, "PPM:H3", 0.9)


# ---- tracer_hordiff's iteration (MOM_tracer_hor_diff.F90:539-607) on two GPUs: tracers-only group pass per iteration ----
def _worker_hordiff(rank, world, port, layout, q, pull):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from mom6_b200 import synth, decomp, multi, capi, tracer_hor_diff
        from mom6_b200.tracer_advect import tracer_advect_init, tracer_advect_end
        from oracle import oracle
        c = synth.make_case(96, 70, 3, 5, cfl=0.5)
        kx, ky = synth.make_khdt(c, 0.3)
        nit = tracer_hor_diff.num_itts_from_cfl(kx, ky, c.G)
        params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": "PPM"}
        h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
        CSo = tracer_advect_init(None, c.G, None, params, None, None)
        Ro = synth.bind_registry(c)
        for _ in range(2):      # step_tracer: advect_tracer, then tracer_hordiff (MOM.F90:1437-1441)
            oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CSo, Ro)
            oracle.tracer_hordiff(h, kx, ky, nit, c.G, c.GV, Ro)
        pe = multi.rank_to_pe(rank, layout)
        G = decomp.local_grid(c.G, layout, pe)
        CS = tracer_advect_init(None, G, None, params, None, None)
        plan = capi.Plan(G, c.GV, CS)
        plan.set_option("pull", pull)
        multi.comm_init_plan(plan, G)
        plan.set_ntr(c.ntr)
        R0 = synth.bind_registry(c)
        for m in range(c.ntr):
            plan.upload_tracer(m, decomp.cut_h(c.G, G, R0.Tr[m].t))
            plan.bind_tracer(m, None, c.conc_underflow(m))
        plan.upload_field(capi.F_H_END, decomp.cut_h(c.G, G, h))
        plan.upload_field(capi.F_UHTR, decomp.cut_u(c.G, G, u))
        plan.upload_field(capi.F_VHTR, decomp.cut_v(c.G, G, v))
        kxl, kyl = decomp.cut_u(c.G, G, kx), decomp.cut_v(c.G, G, ky)
        for _ in range(2):
            plan.advect_resident(c.dt)
            plan.hordiff(kxl, kyl, nit, c.GV.H_subroundoff)
        ok = True
        for m in range(c.ntr):
            o = np.zeros(G.shape_h(c.GV.ke))
            plan.download_tracer(m, o)
            want = decomp.cut_h(c.G, G, Ro.Tr[m].t)
            ok = ok and np.array_equal(G.interior(want).view(np.uint64), G.interior(o).view(np.uint64))
        plan.destroy()
        q.put((rank, bool(ok), nit, nit))
    except Exception as e:      # report instead of leaving the parent to wait for its queue time-out
        q.put((rank, False, -1, "worker raised %s: %s" % (type(e).__name__, str(e)[:300])))
    finally:
        dist.destroy_process_group()


def _run_hordiff(world, layout, pull=1):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [ctx.Process(target=_worker_hordiff, args=(r, world, port, layout, q, pull)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert [r[:2] for r in res] == [(r, True) for r in range(world)], res


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("layout", [(1, 2), (2, 1)])
@pytest.mark.parametrize("pull", [1, 0])
def test_two_gpu_hordiff_layout_invariance(layout, pull):
    _run_hordiff(2, layout, pull)


# ---- the offline advection loop (MOM_offline_main.F90:254-415) on two GPUs: group passes + integer all-reduce ----
def _worker_offline(rank, world, port, layout, q, pull):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from mom6_b200 import synth, decomp, multi, capi
        from mom6_b200.tracer_advect import tracer_advect_init
        from oracle import oracle
        units = dict(H_to_kg_m2=1035.0, L_to_m=1.0, kg_m2_to_H=1.0 / 1035.0, m_to_L=1.0)
        c = synth.make_case(96, 70, 3, 4, cfl=2.4)
        params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": "PPM"}
        h, u, v = c.h_end.numpy().copy(), c.uhtr.numpy().copy(), c.vhtr.numpy().copy()
        CSo = tracer_advect_init(None, c.G, None, params, None, None)
        Ro = synth.bind_registry(c)
        conv_o, nit_o, res_o = oracle.offline_advection_loop(h, u, v, c.dt, 7, c.G, c.GV, CSo, Ro, x_before_y=True, **units)
        pe = multi.rank_to_pe(rank, layout)
        G = decomp.local_grid(c.G, layout, pe)
        CS = tracer_advect_init(None, G, None, params, None, None)
        plan = capi.Plan(G, c.GV, CS)
        plan.set_option("pull", pull)
        multi.comm_init_plan(plan, G)
        plan.set_ntr(c.ntr)
        R0 = synth.bind_registry(c)
        for m in range(c.ntr):
            plan.upload_tracer(m, decomp.cut_h(c.G, G, R0.Tr[m].t))
            plan.bind_tracer(m, None, c.conc_underflow(m))
        plan.upload_field(capi.F_H_END, decomp.cut_h(c.G, G, c.h_end.numpy()))
        plan.upload_field(capi.F_UHTR, decomp.cut_u(c.G, G, c.uhtr.numpy()))
        plan.upload_field(capi.F_VHTR, decomp.cut_v(c.G, G, c.vhtr.numpy()))
        conv, nit, res = plan.offline_advect(c.dt, 7, x_before_y=True, **units)
        bad = []
        if conv != conv_o or nit != nit_o or nit < 2: bad.append("trips %s/%s" % (nit, nit_o))
        if not np.array_equal(res.view(np.uint64), res_o.view(np.uint64)):
            bad.append("residuals first differ at %d" % int(np.argmax(res.view(np.uint64) != res_o.view(np.uint64))))
        nk = c.GV.ke
        for m in range(c.ntr):
            o = np.zeros(G.shape_h(nk))
            plan.download_tracer(m, o)
            want = decomp.cut_h(c.G, G, Ro.Tr[m].t)
            if not np.array_equal(G.interior(want).view(np.uint64), G.interior(o).view(np.uint64)): bad.append("tracer %d" % m)
        ho, uo, vo = np.zeros(G.shape_h(nk)), np.zeros(G.shape_u(nk)), np.zeros(G.shape_v(nk))
        plan.download_field(capi.F_H_END, ho); plan.download_field(capi.F_UHR, uo); plan.download_field(capi.F_VHR, vo)
        if not np.array_equal(G.interior(decomp.cut_h(c.G, G, h)).view(np.uint64), G.interior(ho).view(np.uint64)): bad.append("h_pre")
        if not np.array_equal(G.interior_u(decomp.cut_u(c.G, G, u)).view(np.uint64), G.interior_u(uo).view(np.uint64)): bad.append("uhtr")
        if not np.array_equal(G.interior_v(decomp.cut_v(c.G, G, v)).view(np.uint64), G.interior_v(vo).view(np.uint64)): bad.append("vhtr")
        plan.destroy()
        q.put((rank, not bad, nit, "; ".join(bad)))
    except Exception as e:      # report instead of leaving the parent to wait for its queue time-out
        q.put((rank, False, -1, "worker raised %s: %s" % (type(e).__name__, str(e)[:300])))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("layout", [(1, 2), (2, 1)])
@pytest.mark.parametrize("pull", [1, 0])
def test_two_gpu_offline_loop_layout_invariance(layout, pull):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [ctx.Process(target=_worker_offline, args=(r, 2, port, layout, q, pull)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert [r[:2] for r in res] == [(r, True) for r in range(2)], res


# ---- open boundaries on a multi-block plan: every rank holds its clipped, locally indexed segments --------------------
def _worker_obc(rank, world, port, layout, scheme, x_first, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from mom6_b200 import synth, decomp, multi
        from mom6_b200.types import TracerRegistry, register_tracer
        from mom6_b200.tracer_advect import tracer_advect_init, tracer_advect_end, advect_tracer
        from oracle import oracle
        c, OBC = synth.make_obc_case()
        params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}
        h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
        nk = c.GV.ke
        # (a) the reference algorithm run per PE on this layout (CPU emulation: oracle per rank + emulated group pass)
        R0 = synth.bind_registry(c)
        ranks = []
        for pj in range(layout[1]):
            for pi in range(layout[0]):
                Gr = decomp.local_grid(c.G, layout, (pi, pj))
                Rr = TracerRegistry()
                for tr in R0.Tr:
                    register_tracer(decomp.cut_h(c.G, Gr, tr.t), Rr, underflow_conc=tr.conc_underflow)
                ranks.append(dict(G=Gr, GV=c.GV, Reg=Rr, h_end=decomp.cut_h(c.G, Gr, h), uhtr=decomp.cut_u(c.G, Gr, u),
                                  vhtr=decomp.cut_v(c.G, Gr, v), OBC=decomp.cut_obc(OBC, c.G, Gr),
                                  uhr_out=np.zeros(Gr.shape_u(nk)), vhr_out=np.zeros(Gr.shape_v(nk))))
        CSo = tracer_advect_init(None, c.G, None, params, None, None)
        nits_o = None
        for _ in range(2):
            nits_o = oracle.advect_tracer_multi(ranks, c.G.Domain, CSo, c.dt, x_first_in=x_first)
        # (b) the 1-PE answer, and the answer without the OBC
        R1, Rn = synth.bind_registry(c), synth.bind_registry(c)
        CS1 = tracer_advect_init(None, c.G, None, params, None, None)
        for _ in range(2):
            oracle.advect_tracer(h, u, v, OBC, c.dt, c.G, c.GV, None, CS1, R1, x_first_in=x_first)
        oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, tracer_advect_init(None, c.G, None, params, None, None), Rn,
                             x_first_in=x_first)
        # (c) this rank on its GPU
        pe = multi.rank_to_pe(rank, layout)
        G = decomp.local_grid(c.G, layout, pe)
        OBCl = decomp.cut_obc(OBC, c.G, G)
        R0 = synth.bind_registry(c)
        Reg = TracerRegistry()
        for tr in R0.Tr:
            register_tracer(decomp.cut_h(c.G, G, tr.t), Reg, underflow_conc=tr.conc_underflow)
        CS = tracer_advect_init(None, G, None, params, None, None)
        CS.gpu_options["obc_multi"] = 1          # the path is refused (status 3) without this option
        hl, ul, vl = decomp.cut_h(c.G, G, h), decomp.cut_u(c.G, G, u), decomp.cut_v(c.G, G, v)
        uo, vo = np.zeros(G.shape_u(nk)), np.zeros(G.shape_v(nk))
        nit = 0
        for _ in range(2):
            nit = advect_tracer(hl, ul, vl, OBCl, c.dt, G, c.GV, None, CS, Reg, x_first_in=x_first, uhr_out=uo, vhr_out=vo)
        me = ranks[rank]
        bad, n1pe = [], 0
        if nit != nits_o[rank]: bad.append("iterations %d/%d" % (nit, nits_o[rank]))
        if not any(s.is_E_or_W or s.is_N_or_S for s in OBCl.segment): bad.append("no segment on this PE")
        for m in range(c.ntr):
            if not np.array_equal(G.interior(me["Reg"].Tr[m].t).view(np.uint64), G.interior(Reg.Tr[m].t).view(np.uint64)):
                bad.append("tracer %d differs from the per-PE reference run" % m)
            n1pe += int((G.interior(decomp.cut_h(c.G, G, R1.Tr[m].t)).view(np.uint64) != G.interior(Reg.Tr[m].t).view(np.uint64)).sum())
        if not np.array_equal(G.interior_u(me["uhr_out"]).view(np.uint64), G.interior_u(uo).view(np.uint64)): bad.append("uhr_out")
        if not np.array_equal(G.interior_v(me["vhr_out"]).view(np.uint64), G.interior_v(vo).view(np.uint64)): bad.append("vhr_out")
        if not any(not np.array_equal(c.G.interior(a.t), c.G.interior(b.t)) for a, b in zip(Rn.Tr, R1.Tr)):
            bad.append("the OBC does not matter in this case")
        tracer_advect_end(CS)
        q.put((rank, not bad, n1pe, "; ".join(bad)))
    except Exception as e:      # report instead of leaving the parent to wait for its queue time-out
        q.put((rank, False, -1, "worker raised %s: %s" % (type(e).__name__, str(e)[:300])))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(_ngpu() < 2, reason="needs 2 GPUs")
@pytest.mark.parametrize("layout", [(1, 2), (2, 1)])
@pytest.mark.parametrize("scheme,x_first", [("PPM", True), ("PLM", True), ("PPM:H3", True), ("PPM", False)])
def test_two_gpu_open_boundaries(layout, scheme, x_first):
    """Open boundaries on a multi-block plan: each GPU equals the reference algorithm run per PE on the same layout
    (oracle per rank + emulated group pass, tests/test_oracle.py), bitwise, tracers and remaining transports.  Against
    the 1-PE answer: equal on the y-split layout and for y-first sweeps; on x-split layouts with x-first sweeps the
    per-PE run itself differs from the 1-PE run at a few cells in this basin, with or without the OBC
    (tests/test_oracle.py::test_obc_basin_layouts) -- the GPUs follow the per-PE run."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    procs = [ctx.Process(target=_worker_obc, args=(r, 2, port, layout, scheme, x_first, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert [(r[0], r[1]) for r in res] == [(r, True) for r in range(2)], res
    if layout == (1, 2) or not x_first:
        assert sum(r[2] for r in res) == 0, res          # and equal to the 1-PE answer
