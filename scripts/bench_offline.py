"""Timing of mom6adv_offline_advect (the iteration loop of offline_advection_ale, MOM_offline_main.F90:254-415) on
resident OM4_025-shaped fields, one GPU: ms per trip through the loop, and the parts of a trip that are not
advect_tracer (h_vol, h_new / uhtr_sub / vhtr_sub, three passes, remaining_transport_sum + its host readback).
Warm-ups first; the plan's CUDA events bracket each call.  Prints one JSON line.

  python scripts/bench_offline.py [--workload OM4_025] [--cfl 2.4] [--iters 6] [--steps 3] [--warmup 2]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="OM4_025")
    ap.add_argument("--cfl", type=float, default=2.4)
    ap.add_argument("--iters", type=int, default=6)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    args = ap.parse_args()
    import numpy as np
    import torch
    import bench
    from mom6_b200 import synth, capi
    from mom6_b200.tracer_advect import tracer_advect_init
    if not torch.cuda.is_available():
        raise SystemExit("needs a CUDA device: there is no CPU fallback")
    ni, nj, nk, ntr, scheme = bench.workload_shape(args.workload)
    N = ni * nj * nk
    c = synth.make_case(ni, nj, nk, ntr, cfl=args.cfl, device="cuda")
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    plan = capi.Plan(c.G, c.GV, CS)
    plan.set_ntr(ntr)
    for m in range(ntr):
        t = c.tracer(m)
        plan.upload_tracer(m, t)
        plan.bind_tracer(m, None, c.conc_underflow(m))
        del t
    units = dict(H_to_kg_m2=1035.0, L_to_m=1.0, kg_m2_to_H=1.0 / 1035.0, m_to_L=1.0)

    def call(iters):
        plan.upload_field(capi.F_H_END, c.h_end); plan.upload_field(capi.F_UHTR, c.uhtr); plan.upload_field(capi.F_VHTR, c.vhtr)
        out = plan.offline_advect(c.dt, iters, **units)
        return out, plan.last_advect_ms

    for _ in range(args.warmup):
        call(args.iters)
    sampler = bench.ClockSampler(0)
    sampler.start()
    l0 = plan.kernel_launches
    per, trips, res = [], None, None
    for _ in range(args.steps):
        (conv, nit, r), ms = call(args.iters)
        per.append(ms); trips = nit; res = r
    launches = (plan.kernel_launches - l0) / args.steps
    clocks = sampler.stop()
    # one more call with the per-class event timing on: what of a call is advect_tracer's kernel classes
    plan.set_option("profile", 1)
    (_, nit_p, _), ms_p = call(args.iters)
    kc = {capi.KC_NAMES[k]: plan.kernel_class_time(k)[0] for k in range(len(capi.KC_NAMES))}
    plan.set_option("profile", 0)
    in_advect = sum(kc.values())
    ms = float(np.median(per))
    out = {"what": "mom6adv_offline_advect", "workload": args.workload, "ni": ni, "nj": nj, "nk": nk, "ntr": ntr, "cfl99": args.cfl,
           "num_off_iter": args.iters, "trips": int(trips), "converged": bool(conv), "ms_per_call": ms,
           "ms_per_call_minmax": [min(per), max(per)], "ms_per_trip": ms / max(trips, 1),
           "profiled_call_ms": ms_p, "profiled_call_advect_tracer_classes_ms": kc,
           "ms_per_trip_outside_advect_tracer_classes": (ms_p - in_advect) / max(nit_p, 1),
           "remaining_transport_sum": [float(x) for x in res],
           "cell_tracer_updates_per_s": N * ntr * trips / (ms * 1e-3), "gpu_launches_per_call": launches,
           "steps": args.steps, "warmup": args.warmup, "clocks": clocks, "dtype": "f64", "l2": "inputs_exceed_l2"}
    plan.destroy()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
