#!/usr/bin/env python
"""bench.py -- advect_tracer cell-tracer updates/s on synthetic MOM6 shapes.

    python bench.py --gpus N --steps K --warmup W            (own arm, CUDA)
    python bench.py --impl reference --gpus N --steps K --warmup W   (CPU arm)

A "step" is one complete advect_tracer call (prologue, group pass, both Strang
sweeps, every residual iteration) on the workload named in config.workload.

Own arm: inputs resident in HBM (RESIDENT mode of the C ABI); `value` is
cell-tracer updates/s over the K timed calls, CUDA-event timed on the plan's
stream.  `e2e` repeats the measurement through the reference-facing call
advect_tracer(h_end, uhtr, vhtr, ...) (MIRROR mode) with pinned HOST arrays:
every step copies (ntr+3) fields host->device and ntr fields device->host
inside the timed region.  `roofline` is measured live with CUDA events around
the dominant kernel class; `cpu_baseline` times the CPU oracle (a port of the
reference's arithmetic: no Fortran compiler exists on the box) on a bounded
sample of the same workload.

Reference arm (--impl reference): the same oracle with every host thread on a
bounded sample (a slab of levels of the same horizontal shape, all tracers).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "advect_tracer cell-tracer updates/s"
UNIT = "cell-tracer updates/s"


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int = 0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def apply_env_options(plan):
    """MOM6ADV_OPTS="key=value,..." -> mom6adv_set_option (debugging / A-B runs; the default run sets nothing)."""
    for kv in os.environ.get("MOM6ADV_OPTS", "").split(","):
        if "=" in kv:
            k, v = kv.split("=", 1)
            plan.set_option(k.strip(), int(v))


def shared_config(name, ni, nj, nk, ntr, scheme, cfl, layout="1x1"):
    """The workload description both arms print: identical keys and values for the same workload."""
    return {"workload": name, "ni": ni, "nj": nj, "nk": nk, "ntr": ntr, "scheme": scheme, "halo": 4, "cfl99": cfl,
            "layout": layout, "l2": "inputs_exceed_l2", "fma_contraction": "off"}


def bitcount_chksum_device(t):
    """MOM's bitcount checksum of a device tensor (src/framework/MOM_checksums.F90:1396-1404,2407-2414):
    sum of popcnt(transfer(abs(x), 1_8)), before the modulus.  Integer tensor ops only (torch has no popcount)."""
    import torch
    x = t.abs().contiguous().view(torch.int64)
    x = x - ((x >> 1) & 0x5555555555555555)
    x = (x & 0x3333333333333333) + ((x >> 2) & 0x3333333333333333)
    x = (x + (x >> 4)) & 0x0F0F0F0F0F0F0F0F
    x = (x * 0x0101010101010101) >> 56
    return int((x & 0xFF).sum().item())


BC_MODULUS = 1000000000      # MOM_checksums.F90:95


def tracer_checksums(plan, G, nk, ntr):
    """Per-tracer bitcount sums over the computational domain of the resident tracers (before the modulus, so that
    the sums of the ranks of a multi-GPU run add up to the 1-GPU value)."""
    import torch
    buf = torch.empty(G.shape_h(nk), dtype=torch.float64, device="cuda")
    out = []
    for m in range(ntr):
        plan.download_tracer(m, buf)
        out.append(bitcount_chksum_device(buf[:, G.pj(G.jsc):G.pj(G.jec) + 1, G.pi(G.isc):G.pi(G.iec) + 1]))
    del buf
    return out


def fold_checksums(sums):
    """One number for the JSON line: MOM's per-field checksums (mod bc_modulus) folded position-dependently."""
    acc = 0
    for m, v in enumerate(sums):
        acc = (acc * 31 + (v % BC_MODULUS) + m) % (1 << 61)
    return acc


def workload_shape(name: str):
    from mom6_b200 import synth
    if "," in name:          # ad-hoc "ni,nj,nk,ntr[,scheme]" (profiling runs)
        f = name.split(",")
        return int(f[0]), int(f[1]), int(f[2]), int(f[3]), (f[4] if len(f) > 4 else "PPM")
    ni, nj, nk, ntr, scheme = synth.CONFIGS[name]
    return ni, nj, nk, ntr, scheme


# ---------------------------------------------------------------------------
# CPU arm (oracle): used for cpu_baseline and for --impl reference
# ---------------------------------------------------------------------------
def cpu_sample_setup(name: str, cfl: float, nk_sample: int):
    import numpy as np
    from mom6_b200 import synth
    from mom6_b200.tracer_advect import tracer_advect_init
    ni, nj, nk, ntr, scheme = workload_shape(name)
    # one level per host thread (the oracle threads over k like the reference, :147,247), at least 16, at most nk:
    # every core of the box is used and the sample stays a few seconds per call
    nks = min(nk, nk_sample if nk_sample > 0 else max(16, os.cpu_count() or 1))
    c = synth.make_case(ni, nj, nks, ntr, cfl=cfl)
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    Reg = synth.bind_registry(c)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    return c, CS, Reg, (h, u, v), ni * nj * nks * ntr, nks


def cpu_time_steps(c, CS, Reg, huv, steps: int, warmup: int, threads: int):
    from oracle import oracle
    h, u, v = huv
    for _ in range(warmup):
        oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CS, Reg, nthreads=threads)
    t0 = time.perf_counter()
    nit = 0
    for _ in range(steps):
        nit = oracle.advect_tracer(h, u, v, None, c.dt, c.G, c.GV, None, CS, Reg, nthreads=threads)
    return (time.perf_counter() - t0) / max(steps, 1), nit


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    name = args.workload
    ni, nj, nk, ntr, scheme = workload_shape(name)
    c, CS, Reg, huv, units, nks = cpu_sample_setup(name, args.cfl, args.cpu_levels)
    cores = os.cpu_count() or 1
    threads = min(cores, nks)     # the oracle threads over k like the reference (:147,247)
    sec, nit = cpu_time_steps(c, CS, Reg, huv, args.steps, args.warmup, threads)
    val = units / sec
    sample = "%dx%dx%d-level slab of the %d-level workload, %d tracers, %s, cfl99=%.2f (%d iteration(s) per call), " \
             "oracle (C++ port, OpenMP over k) on %d of the box's %d host cores" % (
                 ni, nj, nks, nk, ntr, scheme, args.cfl, nit, threads, cores)
    out = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True,
           "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
           "config": shared_config(name, ni, nj, nk, ntr, scheme, args.cfl),
           "iterations_per_call": nit, "host_cores": cores,
           "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
           "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))
    return 0


# ---------------------------------------------------------------------------
# own arm
# ---------------------------------------------------------------------------
def run_native(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py (own arm) needs a CUDA device: there is no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from mom6_b200 import synth, capi
    from mom6_b200.tracer_advect import tracer_advect_init, tracer_advect_end, advect_tracer
    from mom6_b200.types import TracerRegistry, register_tracer

    name = args.workload
    ni, nj, nk, ntr, scheme = workload_shape(name)
    N = ni * nj * nk
    units = N * ntr
    peak, peak_src = _peaks()

    if world > 1:
        from mom6_b200 import multi
        return multi.bench_multi(args, name, world, rank, local_rank, peak, peak_src)

    # ---- resident inputs --------------------------------------------------
    c = synth.make_case(ni, nj, nk, ntr, cfl=args.cfl, device="cuda")
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    plan = capi.Plan(c.G, c.GV, CS)
    plan.set_ntr(ntr)
    if args.fast is not None:
        plan.set_option("fast", args.fast)      # A/B runs: 1 = the round-1 streaming kernels
    apply_env_options(plan)
    diag_flag = np.zeros(8)      # RESIDENT mode: a non-NULL pointer only says "associated" (the values stay in HBM)
    for m in range(ntr):
        t = c.tracer(m)
        plan.upload_tracer(m, t)
        if m < args.diags:       # --diags N: ad_x and ad_y on the first N tracers (T_adx, T_ady, S_adx, S_ady of a diag_table)
            plan.bind_tracer(m, None, c.conc_underflow(m), ad_x=diag_flag, ad_y=diag_flag)
        else:
            plan.bind_tracer(m, None, c.conc_underflow(m))
        del t
    plan.upload_field(capi.F_H_END, c.h_end)
    plan.upload_field(capi.F_UHTR, c.uhtr)
    plan.upload_field(capi.F_VHTR, c.vhtr)
    torch.cuda.synchronize()

    stream = torch.cuda.ExternalStream(plan.stream)
    for _ in range(args.warmup):
        nit = plan.advect_resident(c.dt)
    plan.set_option("profile", 1)
    launches0 = plan.kernel_launches
    sampler = ClockSampler(local_rank)
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(stream)
    per_call = []
    for _ in range(args.steps):
        nit = plan.advect_resident(c.dt)
        per_call.append(plan.last_advect_ms)
    e1.record(stream)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    total_ms = e0.elapsed_time(e1)
    ms_step = total_ms / args.steps
    launches = plan.kernel_launches - launches0
    kc = {capi.KC_NAMES[k]: plan.kernel_class_time(k) for k in range(len(capi.KC_NAMES))}
    plan.set_option("profile", 0)
    value = units / (ms_step * 1e-3)

    # ---- roofline of the dominant kernel class -----------------------------
    dom = max(("tracer_x", "tracer_y"), key=lambda k: kc[k][0])
    dom_ms, dom_n = kc[dom]
    dom_avg_ms = dom_ms / max(dom_n, 1)
    # algorithmic bytes of one tracer-sweep launch: every tracer read once and written once,
    # plus the four tracer-independent mass fields of the sweep read once
    alg_launch = (16.0 * ntr + 32.0) * N
    ach = alg_launch / (dom_avg_ms * 1e-3) / 1e9
    b_alg_call = (16.0 * ntr + 24.0) * N
    path_gbs = b_alg_call / (ms_step * 1e-3) / 1e9
    roofline = {"bound": "hbm", "kernel": "k_%s (tracer sweep, all tracers per launch)" % dom,
                "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                "peak_source": peak_src, "avg_launch_ms": dom_avg_ms, "launches_timed": dom_n,
                "alg_bytes_per_launch": alg_launch,
                "path": {"alg_bytes_per_call": b_alg_call, "achieved": path_gbs, "frac": path_gbs / peak,
                         "note": "whole advect_tracer call: B_alg=(16*ntr+24)*N over ms_per_step"},
                "class_ms_per_step": {k: v[0] / args.steps for k, v in kc.items()}}
    tr_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tr_path):
        try:
            roofline["traffic"] = json.load(open(tr_path)).get(name, {}).get(dom)
            roofline["traffic_source"] = "profiles/traffic.json (ncu dram__bytes capture of this kernel at this shape, not re-measured in this run)"
        except Exception:
            pass

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic",
           "config": shared_config(name, ni, nj, nk, ntr, scheme, args.cfl),
           "iterations_per_call": nit, "resident_gb": plan.device_bytes / 1e9, "tracers_with_ad_x_ad_y": int(args.diags),
           "roofline": roofline, "gpu_launches": int(launches), "clocks": clocks,
           "per_call_ms": {"min": min(per_call), "median": sorted(per_call)[len(per_call) // 2], "max": max(per_call)}}
    # ocean-only rate (land included in `value`, the ocean-only count reported alongside)
    mT = np.asarray(c.G.mask2dT)
    n_ocean = float(c.G.interior(mT).sum())
    out["ocean_cell_tracer_updates_per_s"] = n_ocean * nk * ntr / (ms_step * 1e-3)
    out["ocean_fraction"] = n_ocean / (ni * nj)
    # layout-invariance witness: MOM bitcount checksum of every tracer after the timed calls (N = 2/4/8 print the same)
    try:
        sums = tracer_checksums(plan, c.G, nk, ntr)
        out["checksum"] = {"value": fold_checksums(sums), "calls": args.warmup + args.steps,
                           "tracer0_mod_bc": sums[0] % BC_MODULUS, "tracer_last_mod_bc": sums[-1] % BC_MODULUS,
                           "what": "MOM bitcount checksum (MOM_checksums.F90:1396-1404,2407-2414) of each tracer's "
                                   "computational domain after warmup+steps calls, folded over tracers"}
    except Exception as e:
        out["checksum"] = {"error": str(e)[:200]}
    # parity at the benchmarked shape: a slab of the workload's own state, one call, against the CPU oracle, bitwise
    if not args.no_parity:
        try:
            out["parity_slab"] = run_parity_slab(plan, c, ntr, scheme)
        except Exception as e:
            out["parity_slab"] = {"error": str(e)[:300]}

    # ---- e2e, state resident (the integration BASELINE.json's north_star describes: the 50 passive tracers stay
    # in HBM; each step the host hands over h_end, uhtr, vhtr and the two active tracers T, S from pinned memory
    # and takes T, S back) -- reported beside the host-array call, which stays the `e2e` of the contract -----------
    if not args.no_e2e:
        try:
            out["e2e_resident"] = run_e2e_resident(args, plan, c, ni * nj * nk * ntr)
        except Exception as e:      # never lose the main line over the extra leg
            out["e2e_resident"] = {"error": str(e)[:200]}

    # ---- the widened row beside the path: tracer_hordiff's iteration on the same resident state,
    # timed the same way (its own warm-ups, the plan's events around the call's kernels); not part of `value` ----------
    if not args.no_hordiff:
        try:
            out["hordiff"] = run_hordiff_leg(plan, c, ni, nj, nk, ntr, peak)
        except Exception as e:
            out["hordiff"] = {"error": str(e)[:200]}

    # free the resident plan before the e2e / cpu legs
    plan.destroy()
    del plan, c
    torch.cuda.empty_cache()

    # ---- e2e: reference-facing call with pinned host arrays -----------------
    if not args.no_e2e:
        out["e2e"] = run_e2e(args, name)
    # ---- cpu baseline ----------------------------------------------------------
    if not args.no_cpu:
        cc, CSc, Regc, huv, u_s, nks = cpu_sample_setup(name, args.cfl, args.cpu_levels)
        cores = os.cpu_count() or 1
        threads = min(cores, nks)
        sec, nit_c = cpu_time_steps(cc, CSc, Regc, huv, 2, 1, threads)
        out["cpu_baseline"] = {"value": u_s / sec, "unit": UNIT, "cores": threads, "kind": "port", "host_cores": cores,
                               "sample": "%dx%dx%d-level slab of the %d-level workload, %d tracers, %s, 1 warm-up + 2 timed "
                                         "calls of %.1f s, oracle (C++ port, OpenMP over k) on %d of %d host cores"
                                         % (ni, nj, nks, nk, ntr, scheme, sec, threads, cores)}
    print(json.dumps(out))
    return 0


def run_hordiff_leg(plan, c, ni, nj, nk, ntr, peak, calls=3, warmup=2):
    """mom6adv_hordiff (MOM_tracer_hor_diff.F90:539-607) on the resident tracers: ms per call and the roofline of the
    iteration kernel (algorithmic bytes per iteration = (16*ntr + 24) * cells; the single-iteration call isolates the
    coefficient kernel and the two 2-D uploads).  scripts/bench_hordiff.py is the stand-alone form with its parity slab."""
    import numpy as np
    from mom6_b200 import synth, tracer_hor_diff
    kx, ky = synth.make_khdt(c, 0.3)
    nit = tracer_hor_diff.num_itts_from_cfl(kx, ky, c.G)
    hneg = c.GV.H_subroundoff
    for _ in range(warmup):
        plan.hordiff(kx, ky, nit, hneg)
    l0 = plan.kernel_launches
    per, one = [], []
    for _ in range(calls):
        plan.hordiff(kx, ky, nit, hneg)
        per.append(plan.last_advect_ms)
    launches = plan.kernel_launches - l0
    for _ in range(calls):
        plan.hordiff(kx, ky, 1, hneg)
        one.append(plan.last_advect_ms)
    ms, ms1 = float(np.median(per)), float(np.median(one))
    it_ms = (ms - ms1) / (nit - 1) if nit > 1 else ms
    alg = (16.0 * ntr + 24.0) * ni * nj * nk
    return {"what": "mom6adv_hordiff: the diffusion iteration of tracer_hordiff on the resident tracers",
            "num_itts": nit, "ms_per_call": ms, "ms_per_iteration": it_ms, "ms_coefficients_and_uploads": ms1 - it_ms,
            "tracer_cell_updates_per_s": ni * nj * nk * ntr * nit / (ms * 1e-3), "gpu_launches": int(launches),
            "calls": calls, "warmup": warmup,
            "roofline": {"bound": "hbm", "kernel": "k_hd_tile", "achieved": alg / (it_ms * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": alg / (it_ms * 1e-3) / 1e9 / peak, "alg_bytes_per_launch": alg,
                         "traffic": _recorded_traffic("k_hd_tile") if (ni, nj, nk, ntr) == (1440, 1080, 75, 52) else None}}


def _recorded_traffic(kernel, workload="OM4_025"):
    """dram bytes per launch of `kernel` from the committed ncu capture (profiles/traffic.json), or None."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(workload, {}).get(kernel)
    except Exception:
        return None


def _stats_ms(per):
    p = sorted(per)
    return {"min": p[0] * 1e3, "median": p[len(p) // 2] * 1e3, "max": p[-1] * 1e3}


def run_parity_slab(plan, c, ntr, scheme, levels=4):
    """Parity AT the benchmarked shape.  Levels are independent in advect_tracer (every loop is per k; a level that
    converged is skipped, :248-267), so a slab of the workload's own inputs has, level by level, the answer of the full
    run: the tracers are reset to their synthetic initial state, ONE call runs on the GPU at full size, and `levels`
    levels spread over the column are compared bit for bit with the CPU oracle run on exactly those levels (the
    oracle is the checker here, never the thing measured)."""
    import numpy as np
    import torch
    from mom6_b200 import capi, synth
    from mom6_b200.tracer_advect import tracer_advect_init
    from mom6_b200.types import TracerRegistry, register_tracer, VerticalGrid
    from oracle import oracle
    G, nk = c.G, c.GV.ke
    ks = sorted(set(int(round(x)) for x in np.linspace(0, nk - 1, min(levels, nk))))
    kt = torch.as_tensor(ks, device="cuda")
    t0 = time.perf_counter()
    Reg = TracerRegistry()
    for m in range(ntr):
        t = c.tracer(m)
        plan.upload_tracer(m, t)
        register_tracer(np.ascontiguousarray(t.index_select(0, kt).cpu().numpy()), Reg, name="tr%02d" % m,
                        underflow_conc=c.conc_underflow(m))
        del t
    nit_g = plan.advect_resident(c.dt)
    GVs = VerticalGrid(ke=len(ks), Angstrom_H=c.GV.Angstrom_H, H_subroundoff=c.GV.H_subroundoff)
    CS = tracer_advect_init(None, G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    h, u, v = (np.ascontiguousarray(a.index_select(0, kt).cpu().numpy()) for a in (c.h_end, c.uhtr, c.vhtr))
    nit_o = oracle.advect_tracer(h, u, v, None, c.dt, G, GVs, None, CS, Reg, nthreads=len(ks))
    buf = torch.empty(G.shape_h(nk), dtype=torch.float64, device="cuda")
    bad, worst, detail = 0, 0.0, None
    for m in range(ntr):
        plan.download_tracer(m, buf)
        got = np.ascontiguousarray(G.interior(buf.index_select(0, kt).cpu().numpy()))
        want = np.ascontiguousarray(G.interior(Reg.Tr[m].t))
        ne = got.view(np.uint64) != want.view(np.uint64)
        if ne.any():
            bad += 1
            worst = max(worst, float(np.abs(got - want).max()))
            if detail is None:
                w = np.argwhere(ne)
                mT = G.interior(np.asarray(G.mask2dT))
                detail = {"tracer": m, "cells": int(ne.sum()), "per_level": [int(ne[q].sum()) for q in range(len(ks))],
                          "on_land": int((ne & (mT[None] < 0.5)).sum()),
                          "first": [[int(x) for x in w[q]] + [float(got[tuple(w[q])]), float(want[tuple(w[q])])] for q in range(min(4, len(w)))],
                          "j_range": [int(w[:, 1].min()), int(w[:, 1].max())], "i_range": [int(w[:, 2].min()), int(w[:, 2].max())]}
    return {"result": "bitwise" if bad == 0 else "MISMATCH in %d of %d tracers (max abs %g)" % (bad, ntr, worst),
            "levels": ks, "tracers": ntr, "cells_compared": int(G.ni * G.nj * len(ks) * ntr),
            "iterations_gpu_full_column": int(nit_g), "iterations_oracle_slab": int(nit_o), "first_mismatch": detail,
            "oracle": "oracle/advect_ref.cpp (pinned against the translated reference, tests/test_f2c_reference.py)",
            "seconds": time.perf_counter() - t0}


def run_e2e_resident(args, plan, c, units):
    """RESIDENT-mode step through the C ABI: mom6adv_upload_field x3 + mom6adv_upload_tracer(T, S) from pinned host
    arrays, mom6adv_advect_resident, mom6adv_download_tracer(T, S) into pinned host arrays; the passive tracers never
    leave HBM.  Wall clock around whole steps, device synchronised on both sides."""
    import torch
    from mom6_b200 import capi
    def pinned(a):
        h = torch.empty(a.shape, dtype=torch.float64, pin_memory=True)
        h.copy_(a)
        return h
    hh, hu, hv = pinned(c.h_end), pinned(c.uhtr), pinned(c.vhtr)
    ht = [pinned(c.tracer(0)), pinned(c.tracer(1))]
    h2d = sum(x.numel() * 8 for x in (hh, hu, hv)) + sum(x.numel() * 8 for x in ht)
    d2h = sum(x.numel() * 8 for x in ht)

    def step():
        plan.upload_field(capi.F_H_END, hh)
        plan.upload_field(capi.F_UHTR, hu)
        plan.upload_field(capi.F_VHTR, hv)
        for m in (0, 1):
            plan.upload_tracer(m, ht[m])
        plan.advect_resident(c.dt)
        for m in (0, 1):
            plan.download_tracer(m, ht[m])
        plan.sync()
    for _ in range(args.e2e_warmup):
        step()
    torch.cuda.synchronize()
    per = []
    for _ in range(args.e2e_steps):
        t0 = time.perf_counter()
        step()
        torch.cuda.synchronize()
        per.append(time.perf_counter() - t0)
    sec = sum(per) / len(per)
    return {"value": units / sec, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
            "ms_per_step": sec * 1e3, "steps": args.e2e_steps, "warmup": args.e2e_warmup, "ms_per_step_stats": _stats_ms(per),
            "call": "mom6adv_upload_field(h_end,uhtr,vhtr) + mom6adv_upload_tracer(T,S) -> mom6adv_advect_resident -> "
                    "mom6adv_download_tracer(T,S); 50 passive tracers resident in HBM",
            "result_probe": float(ht[0][0, c.G.jsc, c.G.isc])}


def run_e2e(args, name):
    """The call a Fortran host makes: advect_tracer with HOST arrays (MIRROR mode)."""
    import numpy as np
    import torch
    from mom6_b200 import synth
    from mom6_b200.tracer_advect import tracer_advect_init, tracer_advect_end, advect_tracer
    from mom6_b200.types import TracerRegistry, register_tracer
    ni, nj, nk, ntr, scheme = workload_shape(name)
    c = synth.make_case(ni, nj, nk, ntr, cfl=args.cfl, device="cuda")
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    if args.mirror_chunk is not None:
        CS.gpu_options["mirror_chunk"] = args.mirror_chunk
    if args.mirror_ahead is not None:
        CS.gpu_options["mirror_ahead"] = args.mirror_ahead
    if args.trace:
        CS.gpu_options["trace"] = args.trace
    Reg = TracerRegistry()
    keep = []
    for m in range(ntr):
        t = c.tracer(m)
        ht = torch.empty(t.shape, dtype=torch.float64, pin_memory=True)
        ht.copy_(t)
        del t
        keep.append(ht)
        register_tracer(ht.numpy(), Reg, name="tr%02d" % m, underflow_conc=c.conc_underflow(m))
    host = []
    for a in (c.h_end, c.uhtr, c.vhtr):
        ha = torch.empty(a.shape, dtype=torch.float64, pin_memory=True)
        ha.copy_(a)
        host.append(ha)
    G, GV, dt = c.G, c.GV, c.dt
    del c
    torch.cuda.empty_cache()
    h, u, v = (x.numpy() for x in host)
    h2d = sum(x.numel() * 8 for x in host) + sum(x.numel() * 8 for x in keep)
    d2h = sum(x.numel() * 8 for x in keep)
    for _ in range(args.e2e_warmup):
        advect_tracer(h, u, v, None, dt, G, GV, None, CS, Reg)
    torch.cuda.synchronize()
    per = []
    for _ in range(args.e2e_steps):
        t0 = time.perf_counter()
        advect_tracer(h, u, v, None, dt, G, GV, None, CS, Reg)   # returns after the D2H copies completed
        torch.cuda.synchronize()
        per.append(time.perf_counter() - t0)
    sec = sum(per) / len(per)
    chk = float(keep[0].numpy()[0, G.jsc, G.isc])
    tracer_advect_end(CS)
    return {"value": ni * nj * nk * ntr / sec, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "ms_per_step": sec * 1e3, "steps": args.e2e_steps, "warmup": args.e2e_warmup,
            "ms_per_step_stats": _stats_ms(per),
            "call": "mom6_b200.tracer_advect.advect_tracer -> mom6adv_advect (MIRROR mode, pinned host arrays)",
            "result_probe": chk}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="OM4_025")
    ap.add_argument("--cfl", type=float, default=0.45, help="99th percentile upwind CFL of the synthetic transports")
    ap.add_argument("--cpu-levels", type=int, default=0, help="levels in the bounded CPU sample")
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--e2e-warmup", type=int, default=2)
    ap.add_argument("--diags", type=int, default=0, help="associate ad_x and ad_y on the first N tracers (diagnostic-path timing)")
    ap.add_argument("--no-hordiff", action="store_true", help="skip the tracer_hordiff leg (N=1)")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle comparison of a slab of the workload")
    ap.add_argument("--layout", default=None, help="multi-GPU LAYOUT as IxJ (default: strips in j)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-only", action="store_true", help="print only the e2e leg (pipeline tuning)")
    ap.add_argument("--mirror-chunk", type=int, default=None, help="tracers per pipeline stage of the MIRROR call")
    ap.add_argument("--mirror-ahead", type=int, default=None, help="chunks uploaded in front of the running one")
    ap.add_argument("--fast", type=int, default=None, help="library option 'fast' (kernel generation; A/B timing)")
    ap.add_argument("--trace", type=int, default=0, help="library prints the MIRROR pipeline's stream times (2: per chunk)")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "native":
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    if args.e2e_only:
        import torch
        torch.cuda.set_device(0)
        print(json.dumps(run_e2e(args, args.workload)))
        return 0
    return run_native(args)


if __name__ == "__main__":
    sys.exit(main())
