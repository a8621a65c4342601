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
