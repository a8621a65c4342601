#!/usr/bin/env python
"""Generate tests/golden/*.npz: complete inputs AND oracle outputs of small advect_tracer cases.

The reference (MOM6) ships no golden vector for MOM_tracer_advect.F90 and cannot
be built in this image (no Fortran compiler, un-vendored FMS).  The files are
written from the oracle (oracle/advect_ref.cpp); tests/test_f2c_reference.py then
requires the UNMODIFIED reference file, translated mechanically to C++
(oracle/f2c), to reproduce every one of them bit for bit -- so they are the
reference's answers as far as its source defines them, not only the oracle's.

  python tests/golden/make_golden.py      (re)writes every file listed in CASES
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from mom6_b200 import synth                                    # noqa: E402
from mom6_b200.tracer_advect import tracer_advect_init         # noqa: E402
from oracle import oracle                                      # noqa: E402

DIAGS = ("ad_x", "ad_y", "ad2d_x", "ad2d_y", "advection_xy")

# name: (ni, nj, nk, ntr, scheme, cfl, x_first, ncalls, diags, offline, make_case kwargs)
CASES = {
    "tc1_plm":        (10, 8, 8, 3, "PLM", 0.5, None, 2, False, False, {}),
    "dg_plm":         (44, 40, 2, 10, "PLM", 0.9, True, 2, False, False, {}),
    "dg_ppm":         (44, 40, 2, 10, "PPM", 0.9, True, 2, False, False, {}),
    "dg_ppm_yfirst":  (44, 40, 2, 10, "PPM", 0.9, False, 2, False, False, {}),
    "dg_ppmh3":       (44, 40, 2, 10, "PPM:H3", 0.9, True, 2, False, False, {}),
    "dg_ppm_diags":   (44, 40, 2, 4, "PPM", 0.9, None, 1, True, False, {}),
    "dg_ppm_offline": (44, 40, 2, 4, "PPM", 1.5, None, 2, False, True, {}),
    "closed_nonsym":  (37, 29, 3, 3, "PPM", 0.8, None, 1, False, False, {"symmetric": False, "cyclic_x": False}),
    # open boundaries on all four sides (synth.make_obc_case): kwargs marked "obc" build the OBC as well
    "obc_ppm":        (40, 34, 3, 5, "PPM", 0.6, None, 2, True, False, {"obc": True}),
    "obc_plm_yfirst": (40, 34, 3, 5, "PLM", 0.6, False, 1, False, False, {"obc": True}),
}
GRID_FIELDS = ("areaT", "IareaT", "mask2dT", "mask2dCu", "mask2dCv")


def build_case(spec):
    ni, nj, nk, ntr, scheme, cfl, x_first, ncalls, diags, offline, kw = spec
    if kw.get("obc"):
        c, OBC = synth.make_obc_case(ni, nj, nk, ntr, cfl=cfl)
        c.OBC = OBC
    else:
        c = synth.make_case(ni, nj, nk, ntr, cfl=cfl, **kw)
        c.OBC = None
    Reg = synth.bind_registry(c, diags=diags)
    return c, Reg


def obc_reservoirs(OBC):
    """(key, array) of every allocated tres, in segment / entry order."""
    out = []
    if OBC is not None:
        for n, sg in enumerate(OBC.segment):
            for e, t in enumerate(sg.tr_Reg or []):
                if t.tres is not None:
                    out.append(("obc_tres_%d_%d" % (n, e), t.tres))
    return out


def run_oracle(c, Reg, spec, inputs=None):
    """Run the oracle the way the golden file describes; returns the dict of outputs."""
    ni, nj, nk, ntr, scheme, cfl, x_first, ncalls, diags, offline, kw = spec
    G, nk_ = c.G, c.GV.ke
    CS = tracer_advect_init(None, G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    h, u, v = (inputs[k] for k in ("h_end", "uhtr", "vhtr")) if inputs else (c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy())
    uo, vo = np.zeros(G.shape_u(nk_)), np.zeros(G.shape_v(nk_))
    vol = inputs["vol_prev"].copy() if (inputs and offline) else (offline_volume(c, h) if offline else None)
    nits = []
    for call in range(ncalls):
        nits.append(oracle.advect_tracer(h, u, v, c.OBC, c.dt, G, c.GV, None, CS, Reg, x_first_in=x_first,
                                         vol_prev=vol, max_iter_in=2 if offline else None,
                                         update_vol_prev=True if offline else None, uhr_out=uo, vhr_out=vo))
    out = {"n_iter": np.array(nits), "uhr_out": uo, "vhr_out": vo}
    if offline:
        out["vol_prev_out"] = vol
    for m, tr in enumerate(Reg.Tr):
        out["t_out_%02d" % m] = tr.t.copy()
        if diags:
            for d in DIAGS:
                out["%s_%02d" % (d, m)] = getattr(tr, d).copy()
    return out


def offline_volume(c, h):
    G = c.G
    vol = np.zeros(G.shape_h(c.GV.ke))
    G.interior(vol)[...] = G.interior(G.areaT) * G.interior(h) * (1.0 + 0.1 * np.cos(np.arange(G.interior(h).size)).reshape(G.interior(h).shape))
    return vol


def main():
    for name, spec in CASES.items():
        c, Reg = build_case(spec)
        G = c.G
        data = {"h_end": c.h_end.numpy().copy(), "uhtr": c.uhtr.numpy().copy(), "vhtr": c.vhtr.numpy().copy(),
                "dt": np.array(c.dt), "dt_baroclinic": np.array(c.dt_baroclinic),
                "Angstrom_H": np.array(c.GV.Angstrom_H), "H_subroundoff": np.array(c.GV.H_subroundoff),
                "conc_underflow": np.array([c.conc_underflow(m) for m in range(c.ntr)])}
        for f in GRID_FIELDS:
            data["G_" + f] = getattr(G, f).copy()
        for m, tr in enumerate(Reg.Tr):
            data["t_in_%02d" % m] = tr.t.copy()
        if spec[9]:
            data["vol_prev"] = offline_volume(c, data["h_end"])
        for key, a in obc_reservoirs(c.OBC):
            data[key] = a.copy()
        data.update(run_oracle(c, Reg, spec))
        path = os.path.join(HERE, name + ".npz")
        np.savez_compressed(path, **data)
        print("%-16s %7.1f KB  n_iter %s" % (name, os.path.getsize(path) / 1e3, data["n_iter"].tolist()))


if __name__ == "__main__":
    main()
