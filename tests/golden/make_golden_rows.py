#!/usr/bin/env python
"""Generate tests/golden/row_*.npz: answers of the two rows beside advect_tracer, written by the REFERENCE'S OWN LINES.

  hordiff_*  : MOM_tracer_hor_diff.F90:539-607 (the diffusion iteration of tracer_hordiff)
  offline_*  : MOM_offline_main.F90:270-415 with remaining_transport_sum (:634-648), around the translated advect_tracer

Both are the unmodified reference statements translated mechanically to C++ (oracle/f2c, bounds-checked build) and run
here, where /root/reference exists; the files let the oracle and the CUDA path be checked against them where it does
not.  Inputs are the deterministic synthetic cases of mom6_b200.synth (their bit patterns are fingerprinted in the file).

  python tests/golden/make_golden_rows.py      (re)writes every file listed below
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from mom6_b200 import synth, tracer_hor_diff                   # noqa: E402
from mom6_b200.tracer_advect import tracer_advect_init         # noqa: E402
from oracle import f2c_ref, oracle                             # noqa: E402

UNITS = dict(H_to_kg_m2=1035.0, L_to_m=1.0, kg_m2_to_H=1.0 / 1035.0, m_to_L=1.0)

# name: (ni, nj, nk, ntr, make_case kwargs, (Diffuse_ML_interior, nkml, nk_rho_varies, ML_KhTr_scale))
HORDIFF = {
    "hordiff_cyclic":   (28, 22, 3, 3, {}, (False, 0, 0, 0.0)),
    "hordiff_tripolar_ml": (28, 22, 4, 3, {"tripolar_n": True}, (True, 1, 2, 0.5)),
    "hordiff_closed_nonsym": (27, 21, 2, 2, {"cyclic_x": False, "symmetric": False}, (False, 0, 0, 0.0)),
}
# name: (ni, nj, nk, ntr, scheme, cfl, x_before_y, num_off_iter, min_residual, make_case kwargs)
OFFLINE = {
    "offline_ppm":        (30, 24, 2, 3, "PPM", 2.4, True, 8, 0.0, {}),
    "offline_plm_closed": (30, 24, 2, 3, "PLM", 2.4, False, 8, 0.0, {"cyclic_x": False}),
    "offline_ppm_cap":    (30, 24, 2, 3, "PPM:H3", 2.4, True, 2, 0.0, {"cyclic_y": True}),
}


def fingerprint(*arrays) -> np.ndarray:
    """Order-dependent 64-bit fold of the bit patterns of the inputs (detects any drift of the synthetic generator)."""
    acc = np.uint64(1469598103934665603)
    for a in arrays:
        b = np.ascontiguousarray(a, dtype=np.float64).view(np.uint64).ravel()
        with np.errstate(over="ignore"):
            acc = acc * np.uint64(1099511628211) + np.bitwise_xor.reduce(b * (np.arange(b.size, dtype=np.uint64) * np.uint64(2) + np.uint64(1)))
    return np.array([acc], dtype=np.uint64)


def hordiff_case(spec):
    ni, nj, nk, ntr, kw, ml = spec
    c = synth.make_case(ni, nj, nk, ntr, cfl=0.4, **kw)
    kx, ky = synth.make_khdt(c, 0.3)
    return c, kx, ky, tracer_hor_diff.num_itts_from_cfl(kx, ky, c.G)


def run_hordiff(fn, spec, with_dt):
    c, kx, ky, nit = hordiff_case(spec)
    ml = spec[5]
    Reg = synth.bind_registry(c)
    args = dict(Diffuse_ML_interior=ml[0], nkml=ml[1], nk_rho_varies=ml[2], ML_KhTr_scale=ml[3])
    if with_dt:
        fn(c.h_end.numpy(), kx, ky, nit, c.dt, c.G, c.GV, Reg, **args)
    else:
        fn(c.h_end.numpy(), kx, ky, nit, c.G, c.GV, Reg, **args)
    out = {"num_itts": np.array([nit]), "inputs": fingerprint(c.h_end.numpy(), kx, ky, *[synth.bind_registry(c).Tr[m].t for m in range(c.ntr)])}
    for m in range(c.ntr):
        out["t_%d" % m] = Reg.Tr[m].t.copy()
    return out


def offline_case(spec):
    ni, nj, nk, ntr, scheme, cfl, xby, nit, minres, kw = spec
    c = synth.make_case(ni, nj, nk, ntr, cfl=cfl, **kw)
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    return c, CS


def run_offline(fn, spec):
    ni, nj, nk, ntr, scheme, cfl, xby, nit, minres, kw = spec
    c, CS = offline_case(spec)
    Reg = synth.bind_registry(c)
    h, u, v = c.h_end.numpy().copy(), c.uhtr.numpy().copy(), c.vhtr.numpy().copy()
    conv, trips, res = fn(h, u, v, c.dt, nit, c.G, c.GV, CS, Reg, x_before_y=xby, min_residual=minres, **UNITS)
    out = {"converged": np.array([int(conv)]), "trips": np.array([trips]), "residuals": res, "h_pre": h, "uhtr": u, "vhtr": v,
           "inputs": fingerprint(c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy(), *[synth.bind_registry(c).Tr[m].t for m in range(c.ntr)])}
    for m in range(c.ntr):
        out["t_%d" % m] = Reg.Tr[m].t.copy()
    return out


def main():
    if not f2c_ref.available():
        raise SystemExit("the translated reference (oracle/_ref, built from /root/reference) is needed to write these files")
    for name, spec in HORDIFF.items():
        np.savez_compressed(os.path.join(HERE, "row_%s.npz" % name), **run_hordiff(f2c_ref.tracer_hordiff, spec, True))
        print("wrote row_%s.npz" % name)
    for name, spec in OFFLINE.items():
        np.savez_compressed(os.path.join(HERE, "row_%s.npz" % name), **run_offline(f2c_ref.offline_advection_loop, spec))
        print("wrote row_%s.npz" % name)


if __name__ == "__main__":
    main()
