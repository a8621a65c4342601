"""Committed golden vectors (tests/golden/*.npz: full inputs + expected outputs).

CPU: the oracle must reproduce the stored answer bit for bit (pins the oracle across rebuilds).
GPU: the CUDA path, through the C ABI, must reproduce the same stored answer.
Nothing here regenerates inputs: grid metrics, thicknesses, transports and tracers all come from the file."""
import importlib.util
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
mg = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(mg)


def _load(name):
    spec = mg.CASES[name]
    z = dict(np.load(os.path.join(HERE, "golden", name + ".npz")))
    c, Reg = mg.build_case(spec)          # shapes and index ranges only; every array is overwritten from the file
    for f in mg.GRID_FIELDS:
        getattr(c.G, f)[...] = z["G_" + f]
    assert float(z["dt"]) == c.dt and float(z["dt_baroclinic"]) == c.dt_baroclinic
    assert float(z["Angstrom_H"]) == c.GV.Angstrom_H and float(z["H_subroundoff"]) == c.GV.H_subroundoff
    for m, tr in enumerate(Reg.Tr):
        tr.t[...] = z["t_in_%02d" % m]
        assert tr.conc_underflow == z["conc_underflow"][m]
    for key, a in mg.obc_reservoirs(c.OBC):
        a[...] = z[key]
    return spec, z, c, Reg


def _check(c, Reg, spec, z, got):
    G = c.G
    eq = lambda a, b: np.array_equal(np.ascontiguousarray(a).view(np.uint64), np.ascontiguousarray(b).view(np.uint64))
    assert got["n_iter"].tolist() == z["n_iter"].tolist()
    for m in range(len(Reg.Tr)):
        assert eq(G.interior(got["t_out_%02d" % m]), G.interior(z["t_out_%02d" % m])), "tracer %d" % m
        if spec[8]:
            for d in mg.DIAGS:
                a, b = got["%s_%02d" % (d, m)], z["%s_%02d" % (d, m)]
                cut = G.interior_u if d in ("ad_x", "ad2d_x") else G.interior_v if d in ("ad_y", "ad2d_y") else G.interior
                assert eq(cut(a), cut(b)), "%s of tracer %d" % (d, m)
    assert eq(G.interior_u(got["uhr_out"]), G.interior_u(z["uhr_out"]))
    assert eq(G.interior_v(got["vhr_out"]), G.interior_v(z["vhr_out"]))
    if spec[9]:
        assert eq(G.interior(got["vol_prev_out"]), G.interior(z["vol_prev_out"]))


@pytest.mark.parametrize("name", sorted(mg.CASES))
def test_oracle_reproduces_golden(name):
    spec, z, c, Reg = _load(name)
    got = mg.run_oracle(c, Reg, spec, inputs=z)
    _check(c, Reg, spec, z, got)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(mg.CASES))
def test_cuda_reproduces_golden(name):
    from mom6_b200.tracer_advect import tracer_advect_init, tracer_advect_end, advect_tracer
    spec, z, c, Reg = _load(name)
    ni, nj, nk, ntr, scheme, cfl, x_first, ncalls, diags, offline, kw = spec
    G, nk_ = c.G, c.GV.ke
    CS = tracer_advect_init(None, G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}, None, None)
    uo, vo = np.zeros(G.shape_u(nk_)), np.zeros(G.shape_v(nk_))
    vol = z["vol_prev"].copy() if offline else None
    nits = []
    for call in range(ncalls):
        nits.append(advect_tracer(z["h_end"], z["uhtr"], z["vhtr"], c.OBC, c.dt, G, c.GV, None, CS, Reg, x_first_in=x_first,
                                  vol_prev=vol, max_iter_in=2 if offline else None,
                                  update_vol_prev=True if offline else None, uhr_out=uo, vhr_out=vo))
    tracer_advect_end(CS)
    got = {"n_iter": np.array(nits), "uhr_out": uo, "vhr_out": vo, "vol_prev_out": vol}
    for m, tr in enumerate(Reg.Tr):
        got["t_out_%02d" % m] = tr.t
        if diags:
            for d in mg.DIAGS:
                got["%s_%02d" % (d, m)] = getattr(tr, d)
    _check(c, Reg, spec, z, got)
