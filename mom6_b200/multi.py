"""One rank per GPU: 2-D block decomposition of the horizontal domain (MOM6
LAYOUT = npi,npj) with device-resident halo exchanges.

The reference reaches its neighbours through MOM_domains
(create_group_pass / do_group_pass, MOM_tracer_advect.F90:140-144,206, and
sum_across_PEs, :305); here the C library packs the strips on the GPU and moves
them with NCCL send/recv over NVLink (mom6adv_comm_init).  torch.distributed is
only the bootstrap: it carries the 128-byte NCCL id to every rank, exactly what
an MPI_Bcast does in the Fortran model.
"""
from __future__ import annotations

import os
from typing import Sequence, Tuple

import numpy as np

from . import decomp


def choose_layout(nranks: int, nig: int, njg: int) -> Tuple[int, int]:
    """LAYOUT for nranks GPUs: strips in j while rows stay long (cyclic x and
    the E/W halos stay GPU-local, N/S strips are contiguous), else the most
    square factorisation."""
    if nranks <= 4 or njg // nranks >= 64:
        return (1, nranks)
    best = (1, nranks)
    for pi in range(1, nranks + 1):
        if nranks % pi == 0:
            pj = nranks // pi
            if abs(nig / pi - njg / pj) < abs(nig / best[0] - njg / best[1]):
                best = (pi, pj)
    return best


def layout_extents(nig: int, njg: int, layout: Sequence[int]):
    """Global compute extents of every layout column / row (mpp_get_compute_domains)."""
    ei = decomp.compute_extent(nig, layout[0])
    ej = decomp.compute_extent(njg, layout[1])
    return ([a for a, _ in ei], [b for _, b in ei], [a for a, _ in ej], [b for _, b in ej])


def rank_to_pe(rank: int, layout: Sequence[int]) -> Tuple[int, int]:
    return (rank % layout[0], rank // layout[0])


def broadcast_unique_id(rank: int):
    """Rank 0 draws the NCCL id; torch.distributed (any backend) carries it."""
    import torch
    import torch.distributed as dist
    from . import capi
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        t.copy_(torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(t, src=0)
    return bytes(t.cpu().numpy().tobytes())


def comm_init_plan(plan, G) -> None:
    """Give a multi-block plan its exchange lists and its NCCL communicator."""
    import torch.distributed as dist
    D = G.Domain
    if D.layout[0] * D.layout[1] == 1:
        return
    if not dist.is_initialized():
        raise RuntimeError("a multi-block grid needs torch.distributed to be initialised (one rank per GPU)")
    rank = D.pe[0] + D.layout[0] * D.pe[1]
    if rank != dist.get_rank() or dist.get_world_size() != D.layout[0] * D.layout[1]:
        raise RuntimeError("LAYOUT %s / PE %s disagrees with the process group" % (D.layout, D.pe))
    plan.comm_init(broadcast_unique_id(rank), layout_extents(D.nig, D.njg, D.layout))


def cut_local(c_global, layout, pe):
    """Local grid + h_end/uhtr/vhtr blocks of a globally generated SynthCase (torch tensors stay on their device)."""
    import torch
    Gg = c_global.G
    G = decomp.local_grid(Gg, layout, pe)

    def cut(fn, a):
        if isinstance(a, torch.Tensor):
            idx = fn(Gg, G, None, index_only=True)
            rj = torch.as_tensor(idx[0], device=a.device)
            ci = torch.as_tensor(idx[1], device=a.device)
            return a.index_select(-2, rj).index_select(-1, ci).contiguous()
        return fn(Gg, G, a)
    return G, cut(decomp.cut_h, c_global.h_end), cut(decomp.cut_u, c_global.uhtr), cut(decomp.cut_v, c_global.vhtr), cut


def plan_halo_bytes(plan) -> int:
    try:
        return int(plan.halo_bytes)
    except Exception:
        return 0


def bench_multi(args, name, world, rank, local_rank, peak, peak_src):
    """bench.py body for N > 1: the named global grid split over N GPUs (strong scaling)."""
    import json
    import torch
    import torch.distributed as dist
    from . import synth, capi
    from .tracer_advect import tracer_advect_init
    import bench as B
    ni, nj, nk, ntr, scheme = B.workload_shape(name)
    layout = choose_layout(world, ni, nj)
    if getattr(args, "layout", None):      # --layout IxJ: a 2-D decomposition (MOM6's LAYOUT, src/framework/MOM_domains.F90:66)
        layout = tuple(int(x) for x in args.layout.lower().split("x"))
        if layout[0] * layout[1] != world:
            raise SystemExit("--layout %s does not match %d ranks" % (args.layout, world))
    pe = rank_to_pe(rank, layout)
    c = synth.make_case(ni, nj, nk, ntr, cfl=args.cfl, device="cuda")
    G, h, u, v, cut = cut_local(c, layout, pe)
    CS = tracer_advect_init(None, G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    plan = capi.Plan(G, c.GV, CS)
    plan.set_option("pull", int(os.environ.get("MOM6ADV_PULL", "1")))     # 0: NCCL send/recv exchange (comparison runs)
    plan.set_option("overlap", int(os.environ.get("MOM6ADV_OVERLAP", "1")))
    plan.set_option("dirty", int(os.environ.get("MOM6ADV_DIRTY", "1")))
    plan.set_option("pull2", int(os.environ.get("MOM6ADV_PULL2", "1")))
    B.apply_env_options(plan)
    comm_init_plan(plan, G)
    plan.set_ntr(ntr)
    for m in range(ntr):
        t = cut(decomp.cut_h, c.tracer(m))
        plan.upload_tracer(m, t)
        plan.bind_tracer(m, None, c.conc_underflow(m))
        del t
    plan.upload_field(capi.F_H_END, h); plan.upload_field(capi.F_UHTR, u); plan.upload_field(capi.F_VHTR, v)
    dt = c.dt
    del c, h, u, v
    torch.cuda.empty_cache()
    stream = torch.cuda.ExternalStream(plan.stream)
    for _ in range(args.warmup):
        nit = plan.advect_resident(dt)
    plan.set_option("profile", 1)
    l0 = plan.kernel_launches
    sampler = B.ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0.record(stream)
    for _ in range(args.steps):
        nit = plan.advect_resident(dt)
    e1.record(stream)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    clocks = sampler.stop()
    ms = torch.tensor([e0.elapsed_time(e1) / args.steps], device="cuda", dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)            # max over ranks, timed on the device
    ms_step = float(ms.item())
    kc = {capi.KC_NAMES[k]: plan.kernel_class_time(k) for k in range(len(capi.KC_NAMES))}
    launches = plan.kernel_launches - l0
    # per-rank class times (strips carry different amounts of land: the slowest rank sets the pace)
    kc_t = torch.tensor([plan.kernel_class_time(k)[0] / args.steps for k in range(len(capi.KC_NAMES))] + [e0.elapsed_time(e1) / args.steps],
                        device="cuda", dtype=torch.float64)
    kc_all = [torch.zeros_like(kc_t) for _ in range(world)]
    dist.all_gather(kc_all, kc_t)
    by_rank = {capi.KC_NAMES[k]: [round(float(t[k]), 3) for t in kc_all] for k in range(len(capi.KC_NAMES))
               if any(float(t[k]) > 0 for t in kc_all)}
    by_rank["total"] = [round(float(t[-1]), 3) for t in kc_all]
    N = ni * nj * nk
    Nl = G.ni * G.nj * nk
    units = N * ntr
    dom = max(("tracer_x", "tracer_y"), key=lambda k: kc[k][0])
    dom_avg = kc[dom][0] / max(kc[dom][1], 1)
    alg_launch = (16.0 * ntr + 32.0) * Nl
    ach = alg_launch / (dom_avg * 1e-3) / 1e9
    b_alg = (16.0 * ntr + 24.0) * N
    halo_ms = kc["halo"][0] / args.steps
    # layout-invariance witness: the per-tracer bitcount sums of all ranks add up to the 1-GPU sums
    try:
        sums = torch.tensor(B.tracer_checksums(plan, G, nk, ntr), device="cuda", dtype=torch.int64)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
        sums = [int(x) for x in sums.tolist()]
        checksum = {"value": B.fold_checksums(sums), "calls": args.warmup + args.steps,
                    "tracer0_mod_bc": sums[0] % B.BC_MODULUS, "tracer_last_mod_bc": sums[-1] % B.BC_MODULUS,
                    "what": "MOM bitcount checksum of each tracer's computational domain summed over ranks "
                            "(sum_across_PEs, MOM_checksums.F90:1396-1404), folded over tracers: equals the 1-GPU value"}
    except Exception as e:
        checksum = {"error": str(e)[:200]}
    halo_bytes = torch.tensor([float(plan_halo_bytes(plan)) / max(args.steps + args.warmup, 1)], device="cuda", dtype=torch.float64)
    dist.all_reduce(halo_bytes, op=dist.ReduceOp.MAX)
    if rank == 0:
        out = {"metric": B.METRIC, "value": units / (ms_step * 1e-3), "unit": B.UNIT, "n_gpus": world,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
               "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
               "config": B.shared_config(name, ni, nj, nk, ntr, scheme, args.cfl, layout="%dx%d" % tuple(layout)),
               "iterations_per_call": nit, "resident_gb_per_gpu": plan.device_bytes / 1e9, "checksum": checksum,
               "halo_exchange": {"how": ("peer-memory loads over NVLink (CUDA IPC), all-reduce fences" if int(os.environ.get("MOM6ADV_PULL", "1")) else "GPU pack -> NCCL send/recv (NVLink) -> GPU unpack"),
                                 "bytes_pulled_per_call_max_rank": float(halo_bytes.item()),
                                 "nvlink_frac_of_900GBs_if_fully_exposed": (float(halo_bytes.item()) / 900e9) / max(halo_ms * 1e-3, 1e-9)},
               "roofline": {"bound": "hbm", "kernel": "k_%s (rank 0 block)" % dom, "achieved": ach, "peak": peak,
                            "unit": "GB/s", "frac": ach / peak, "traffic": None, "peak_source": peak_src,
                            "avg_launch_ms": dom_avg, "alg_bytes_per_launch": alg_launch,
                            "path": {"alg_bytes_per_call": b_alg, "achieved": b_alg / (ms_step * 1e-3) / 1e9,
                                     "frac": b_alg / (ms_step * 1e-3) / 1e9 / (peak * world),
                                     "note": "whole call, all GPUs: B_alg over ms_per_step over N x peak"},
                            "class_ms_per_step": {k: v[0] / args.steps for k, v in kc.items()},
                            "class_ms_per_step_by_rank": by_rank,
                            "halo_ms_per_step": halo_ms},
               "gpu_launches": int(launches), "clocks": clocks,
               "e2e": None, "cpu_baseline": None}
        print(json.dumps(out))
    plan.destroy()
    dist.barrier()
    dist.destroy_process_group()
    return 0
