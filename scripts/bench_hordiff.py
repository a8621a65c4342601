"""Timing of mom6adv_hordiff (the diffusion iteration of tracer_hordiff, MOM_tracer_hor_diff.F90:539-607) on resident
OM4_025-shaped fields, one GPU.  Not bench.py's metric: a measurement of the widened row to the
same rules -- warm-ups, CUDA events on the plan's stream (plan.last_advect_ms brackets the call's kernels), inputs far
larger than L2, a parity slab against the oracle on the same box.

  python scripts/bench_hordiff.py [--workload OM4_025] [--steps 5] [--warmup 3] [--itts 3]
Prints one JSON line; `roofline.achieved` = algorithmic bytes of the iteration kernel / its time:
(16 B x ntr + 24 B) per cell and iteration (every tracer read and written once, three coefficient fields read)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="OM4_025")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--itts", type=int, default=0, help="num_itts (0: from the diffusive CFL of the synthetic khdt, :368-387)")
    ap.add_argument("--strength", type=float, default=0.3)
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--variants", default="", help="experiment: comma list of MOM6ADV_HD_VARIANT values to time")
    args = ap.parse_args()
    import numpy as np
    import torch
    import bench
    from mom6_b200 import synth, capi, tracer_hor_diff
    from mom6_b200.tracer_advect import tracer_advect_init
    if not torch.cuda.is_available():
        raise SystemExit("needs a CUDA device: there is no CPU fallback")
    ni, nj, nk, ntr, scheme = bench.workload_shape(args.workload)
    N = ni * nj * nk
    peak, peak_src = bench._peaks()
    c = synth.make_case(ni, nj, nk, ntr, cfl=0.3, device="cuda")
    kx, ky = synth.make_khdt(c, args.strength)
    nit = args.itts or tracer_hor_diff.num_itts_from_cfl(kx, ky, c.G)
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    plan = capi.Plan(c.G, c.GV, CS)
    plan.set_ntr(ntr)
    for m in range(ntr):
        t = c.tracer(m)
        plan.upload_tracer(m, t)
        plan.bind_tracer(m, None, c.conc_underflow(m))
        del t
    plan.upload_field(capi.F_H_END, c.h_end)
    torch.cuda.synchronize()
    hneg = c.GV.H_subroundoff
    for _ in range(args.warmup):
        plan.hordiff(kx, ky, nit, hneg)
    if args.variants:
        res = {}
        for v in args.variants.split(","):
            os.environ["MOM6ADV_HD_VARIANT"] = v
            plan.hordiff(kx, ky, nit, hneg)
            tt = []
            for _ in range(3):
                plan.hordiff(kx, ky, nit, hneg)
                tt.append(plan.last_advect_ms)
            res[v] = float(np.median(tt))
        print(json.dumps({"variants_ms_per_call": res, "num_itts": nit}))
        plan.destroy()
        return
    sampler = bench.ClockSampler(0)
    sampler.start()
    l0 = plan.kernel_launches
    per = []
    for _ in range(args.steps):
        plan.hordiff(kx, ky, nit, hneg)
        per.append(plan.last_advect_ms)
    clocks = sampler.stop()
    launches = plan.kernel_launches - l0
    # one more call with a single iteration: the difference isolates the iteration kernel from the coefficient kernel
    one = []
    for _ in range(max(2, args.steps // 2)):
        plan.hordiff(kx, ky, 1, hneg)
        one.append(plan.last_advect_ms)
    ms = float(np.median(per)); ms1 = float(np.median(one))
    it_ms = (ms - ms1) / (nit - 1) if nit > 1 else ms
    alg_iter = (16.0 * ntr + 24.0) * N
    out = {"what": "mom6adv_hordiff", "workload": args.workload, "ni": ni, "nj": nj, "nk": nk, "ntr": ntr,
           "num_itts": nit, "ms_per_call": ms, "ms_per_call_minmax": [min(per), max(per)], "ms_one_iteration_call": ms1,
           "ms_per_iteration_kernel": it_ms, "ms_coefficients_and_uploads": ms1 - it_ms,
           "tracer_cell_updates_per_s": N * ntr * nit / (ms * 1e-3),
           "roofline": {"bound": "hbm", "achieved": alg_iter / (it_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                        "frac": alg_iter / (it_ms * 1e-3) / 1e9 / peak, "peak_source": peak_src, "traffic": None,
                        "algorithmic_bytes_per_launch": alg_iter},
           "gpu_launches": launches, "steps": args.steps, "warmup": args.warmup, "clocks": clocks,
           "l2": "inputs_exceed_l2", "dtype": "f64"}
    if not args.no_parity:
        from oracle import oracle      # the checker, for the slab only
        # parity on the same box: the first levels of the same fields through the oracle (levels are independent)
        lev = 2
        cs = synth.make_case(ni, nj, lev, min(ntr, 3), cfl=0.3)
        ksx, ksy = synth.make_khdt(cs, args.strength)
        Ro = synth.bind_registry(cs)
        oracle.tracer_hordiff(cs.h_end.numpy(), ksx, ksy, nit, cs.G, cs.GV, Ro)
        CS2 = tracer_advect_init(None, cs.G, None, {"DT": cs.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
        p2 = capi.Plan(cs.G, cs.GV, CS2)
        p2.set_ntr(cs.ntr)
        R0 = synth.bind_registry(cs)
        for m in range(cs.ntr):
            p2.upload_tracer(m, R0.Tr[m].t)
            p2.bind_tracer(m, None, cs.conc_underflow(m))
        p2.hordiff(ksx, ksy, nit, cs.GV.H_subroundoff, h=cs.h_end.numpy())
        ok = True
        for m in range(cs.ntr):
            o = np.zeros(cs.G.shape_h(lev))
            p2.download_tracer(m, o)
            ok = ok and np.array_equal(o.view(np.uint64), Ro.Tr[m].t.view(np.uint64))
        out["parity_slab"] = {"levels": lev, "tracers": cs.ntr, "ni": ni, "nj": nj, "bitwise_equal_to_oracle": bool(ok)}
        p2.destroy()
    plan.destroy()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
