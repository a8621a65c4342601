"""TEST INFRASTRUCTURE ONLY -- ctypes wrapper around the CPU oracle
(oracle/advect_ref.cpp, a restatement of
/root/reference/src/tracer/MOM_tracer_advect.F90:52-1087).  Pinned bit for bit against a mechanical translation of
the unmodified reference file (oracle/f2c, tests/test_f2c_reference.py); no gfortran binary has been available.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from typing import List, Optional, Sequence

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle_advect.so")
_lib = None

_dp = C.POINTER(C.c_double)


class ora_grid(C.Structure):
    _fields_ = [(n, C.c_int) for n in
                ("isc", "iec", "jsc", "jec", "isd", "ied", "jsd", "jed",
                 "IsdB", "IedB", "JsdB", "JedB", "nk", "first_direction")] + \
               [(n, _dp) for n in ("areaT", "IareaT", "mask2dT", "mask2dCu", "mask2dCv")] + \
               [("Angstrom_H", C.c_double), ("H_subroundoff", C.c_double)]


class ora_cs(C.Structure):
    _fields_ = [("dt", C.c_double), ("usePPM", C.c_int), ("useHuynh", C.c_int),
                ("useHuynhStencilBug", C.c_int)]


class ora_tracer(C.Structure):
    _fields_ = [(n, _dp) for n in ("t", "ad_x", "ad_y", "ad2d_x", "ad2d_y", "advection_xy")] + \
               [("conc_underflow", C.c_double)]


class ora_domain(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("nig", "njg", "cyclic_x", "cyclic_y", "tripolar_n", "symmetric")]


class ora_rank(C.Structure):
    _fields_ = [("G", ora_grid), ("ioff", C.c_int), ("joff", C.c_int),
                ("Tr", C.POINTER(ora_tracer)),
                ("h_end", _dp), ("uhtr", _dp), ("vhtr", _dp),
                ("vol_prev", _dp), ("uhr_out", _dp), ("vhr_out", _dp),
                ("n_iter", C.c_int)]


class ora_obc_tracer(C.Structure):
    _fields_ = [("ntr_index", C.c_int), ("tres", _dp), ("OBC_inflow_conc", C.c_double)]


class ora_obc_segment(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("is_E_or_W", "is_N_or_S", "direction", "specified",
                                       "isd", "ied", "jsd", "jed", "IsdB", "IedB", "JsdB", "JedB", "ntseg")] + \
               [("Tr", C.POINTER(ora_obc_tracer))]


class ora_obc(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("OBC_pe", "number_of_segments", "open_u_BCs_exist_globally",
                                       "open_v_BCs_exist_globally", "specified_u_BCs_exist_globally",
                                       "specified_v_BCs_exist_globally")] + \
               [("segment", C.POINTER(ora_obc_segment))]


PASS_FN = C.CFUNCTYPE(None, C.c_void_p, _dp, _dp, _dp, C.POINTER(ora_tracer), C.c_int)
SUM_FN = C.CFUNCTYPE(None, C.c_void_p, C.POINTER(C.c_int), C.c_int)


def build(force: bool = False) -> str:
    """Compile the oracle with oracle/Makefile (g++ -O2 -ffp-contract=off -fopenmp)."""
    src = os.path.join(_HERE, "advect_ref.cpp")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _LIB_PATH


def _offline_argtypes():
    return [C.POINTER(ora_grid), C.POINTER(ora_cs), C.POINTER(ora_tracer), C.c_int, _dp, _dp, _dp, C.c_double, C.c_int, C.c_int,
            C.c_int, C.c_double, C.c_double, C.c_double, C.c_double, C.c_double, C.POINTER(ora_domain), C.c_void_p,
            C.POINTER(C.c_int), C.POINTER(C.c_int), _dp, C.c_int]


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.ora_advect_tracer.restype = C.c_int
        _lib.ora_advect_tracer.argtypes = [
            C.POINTER(ora_grid), C.POINTER(ora_cs), C.POINTER(ora_tracer), C.c_int,
            _dp, _dp, _dp, C.c_double, C.c_int, _dp, C.c_int, C.c_int, _dp, _dp,
            C.POINTER(ora_domain), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.c_int]
        _lib.ora_advect_tracer_obc.restype = C.c_int
        _lib.ora_advect_tracer_obc.argtypes = _lib.ora_advect_tracer.argtypes + [C.POINTER(ora_obc)]
        _lib.ora_advect_tracer_multi.restype = C.c_int
        _lib.ora_advect_tracer_multi.argtypes = [
            C.c_int, C.POINTER(ora_rank), C.POINTER(ora_domain), C.POINTER(ora_cs),
            C.c_int, C.c_double, C.c_int, C.c_int, C.c_int]
        _lib.ora_tracer_hordiff.restype = C.c_int
        _lib.ora_tracer_hordiff.argtypes = [C.POINTER(ora_grid), C.POINTER(ora_tracer), C.c_int, _dp, _dp, _dp, C.c_int,
                                            C.c_double, C.c_int, C.c_int, C.c_int, C.c_double, C.POINTER(ora_domain),
                                            C.c_void_p, C.c_void_p]
        _lib.ora_offline_advection_loop.restype = C.c_int
        _lib.ora_offline_advection_loop.argtypes = _offline_argtypes()
        _lib.ora_remaining_transport_sum.restype = C.c_double
        _lib.ora_remaining_transport_sum.argtypes = [C.POINTER(ora_grid), _dp, _dp, _dp, C.c_double, C.c_double, C.c_double, C.c_double]
        _lib.ora_bitcount_chksum.restype = C.c_longlong
        _lib.ora_bitcount_chksum.argtypes = [_dp, C.c_longlong]
        _lib.ora_efp_sum.restype = None
        _lib.ora_efp_sum.argtypes = [_dp, C.c_longlong, C.POINTER(C.c_longlong)]
        _lib.ora_efp_to_real.restype = C.c_double
        _lib.ora_efp_to_real.argtypes = [C.POINTER(C.c_longlong)]
    return _lib


def _p(a: Optional[np.ndarray]):
    if a is None:
        return None
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"], "need contiguous float64"
    return a.ctypes.data_as(_dp)


def make_grid(G, GV) -> ora_grid:
    g = ora_grid()
    for n in ("isc", "iec", "jsc", "jec", "isd", "ied", "jsd", "jed", "IsdB", "IedB", "JsdB", "JedB",
              "first_direction"):
        setattr(g, n, int(getattr(G, n)))
    g.nk = int(GV.ke)
    for n in ("areaT", "IareaT", "mask2dT", "mask2dCu", "mask2dCv"):
        setattr(g, n, _p(getattr(G, n)))
    g.Angstrom_H = GV.Angstrom_H
    g.H_subroundoff = GV.H_subroundoff
    return g


def make_domain(D) -> ora_domain:
    d = ora_domain()
    d.nig, d.njg = D.nig, D.njg
    d.cyclic_x, d.cyclic_y, d.tripolar_n = int(D.cyclic_x), int(D.cyclic_y), int(D.tripolar_n)
    d.symmetric = int(D.symmetric)
    return d


def make_cs(CS) -> ora_cs:
    c = ora_cs()
    c.dt = CS.dt
    c.usePPM, c.useHuynh, c.useHuynhStencilBug = int(CS.usePPM), int(CS.useHuynh), int(CS.useHuynhStencilBug)
    return c


def make_tracers(Reg):
    arr = (ora_tracer * max(Reg.ntr, 1))()
    for m, tr in enumerate(Reg.Tr[:Reg.ntr]):
        arr[m].t = _p(tr.t)
        for n in ("ad_x", "ad_y", "ad2d_x", "ad2d_y", "advection_xy"):
            setattr(arr[m], n, _p(getattr(tr, n)))
        arr[m].conc_underflow = tr.conc_underflow
    return arr


def make_obc(OBC):
    """ocean_OBC_type -> ora_obc (returns the struct and the objects that must stay alive)."""
    if OBC is None:
        return None, []
    keep = []
    segs = (ora_obc_segment * max(OBC.number_of_segments, 1))()
    for n, sg in enumerate(OBC.segment):
        s = segs[n]
        s.is_E_or_W, s.is_N_or_S, s.direction, s.specified = int(sg.is_E_or_W), int(sg.is_N_or_S), int(sg.direction), int(sg.specified)
        for f in ("isd", "ied", "jsd", "jed", "IsdB", "IedB", "JsdB", "JedB"):
            setattr(s, f, int(getattr(sg.HI, f)))
        if sg.tr_Reg is None:
            s.ntseg = -1
        else:
            s.ntseg = len(sg.tr_Reg)
            trs = (ora_obc_tracer * max(len(sg.tr_Reg), 1))()
            for m, t in enumerate(sg.tr_Reg):
                trs[m].ntr_index, trs[m].tres, trs[m].OBC_inflow_conc = int(t.ntr_index), _p(t.tres), float(t.OBC_inflow_conc)
            keep.append(trs)
            s.Tr = C.cast(trs, C.POINTER(ora_obc_tracer))
    o = ora_obc()
    o.OBC_pe, o.number_of_segments = int(OBC.OBC_pe), OBC.number_of_segments
    o.open_u_BCs_exist_globally, o.open_v_BCs_exist_globally = int(OBC.open_u_BCs_exist_globally), int(OBC.open_v_BCs_exist_globally)
    o.specified_u_BCs_exist_globally = int(OBC.specified_u_BCs_exist_globally)
    o.specified_v_BCs_exist_globally = int(OBC.specified_v_BCs_exist_globally)
    o.segment = C.cast(segs, C.POINTER(ora_obc_segment))
    keep.append(segs)
    return o, keep


def _opt_flag(v):
    return -1 if v is None else int(bool(v))


def advect_tracer(h_end, uhtr, vhtr, OBC, dt, G, GV, US, CS, Reg, x_first_in=None,
                  vol_prev=None, max_iter_in=None, update_vol_prev=None, uhr_out=None,
                  vhr_out=None, nthreads: int = 1, pass_fn=None, sum_fn=None) -> int:
    """Oracle advect_tracer with the reference's argument list
    (MOM_tracer_advect.F90:52-53).  Updates Reg.Tr[m].t in place.  Returns the
    number of iterations taken."""
    if CS is None:
        raise RuntimeError("MOM_tracer_advect: tracer_advect_init must be called before advect_tracer.")
    if Reg is None:
        raise RuntimeError("MOM_tracer_advect: register_tracer must be called before advect_tracer.")
    g, cs, dom = make_grid(G, GV), make_cs(CS), make_domain(G.Domain)
    tr = make_tracers(Reg)
    nit = C.c_int(0)
    pf = C.cast(pass_fn, C.c_void_p) if pass_fn is not None else None
    sf = C.cast(sum_fn, C.c_void_p) if sum_fn is not None else None
    obc, _keep = make_obc(OBC)
    rc = lib().ora_advect_tracer_obc(
        C.byref(g), C.byref(cs), tr, Reg.ntr, _p(h_end), _p(uhtr), _p(vhtr), float(dt),
        _opt_flag(x_first_in), _p(vol_prev), -1 if max_iter_in is None else int(max_iter_in),
        _opt_flag(update_vol_prev), _p(uhr_out), _p(vhr_out), C.byref(dom), pf, sf, None,
        C.byref(nit), int(nthreads), C.byref(obc) if obc is not None else None)
    if rc != 0:
        raise RuntimeError("oracle advect_tracer failed: %d" % rc)
    return nit.value


def df_pointer_tables(df, ntr):
    """df = {"df_x": [array or None per tracer], ...} -> four (c_void_p * ntr) tables (None entries == unassociated)."""
    out = []
    for key in ("df_x", "df_y", "df2d_x", "df2d_y"):
        tab = (C.c_void_p * max(ntr, 1))()
        for m, a in enumerate((df or {}).get(key) or []):
            if a is not None:
                assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
                tab[m] = a.ctypes.data
        out.append(tab)
    return out


def tracer_hordiff(h, khdt_x, khdt_y, num_itts, G, GV, Reg, Diffuse_ML_interior=False, nkml=0, nk_rho_varies=0,
                   ML_KhTr_scale=0.0, df=None, dt=None) -> None:
    """The along-layer diffusion iteration of tracer_hordiff (MOM_tracer_hor_diff.F90:539-607) on Reg.Tr[m].t in place;
    khdt_x / khdt_y / num_itts as the reference's 2-D set-up (:186-387) leaves them.  df (optional): the diffusive-flux
    diagnostics of :576-591 as {"df_x": [...], "df_y": [...], "df2d_x": [...], "df2d_y": [...]}, one array or None per
    tracer, accumulated in place (zero them first, :388-405); needs dt (Idt = 1/dt, :204)."""
    g, dom = make_grid(G, GV), make_domain(G.Domain)
    tr = make_tracers(Reg)
    if df is not None:
        tabs = df_pointer_tables(df, Reg.ntr)
        f = lib().ora_tracer_hordiff_df
        f.restype = C.c_int
        rc = f(C.byref(g), tr, C.c_int(Reg.ntr), _p(h), _p(khdt_x), _p(khdt_y), C.c_int(int(num_itts)),
               C.c_double(float(GV.H_subroundoff)), C.c_int(int(bool(Diffuse_ML_interior))), C.c_int(int(nkml)),
               C.c_int(int(nk_rho_varies)), C.c_double(float(ML_KhTr_scale)), C.byref(dom), None, None,
               C.c_double(1.0 / float(dt)), *tabs)
        if rc != 0:
            raise RuntimeError("oracle tracer_hordiff (df) failed: %d" % rc)
        return
    rc = lib().ora_tracer_hordiff(C.byref(g), tr, Reg.ntr, _p(h), _p(khdt_x), _p(khdt_y), int(num_itts),
                                  float(GV.H_subroundoff), int(bool(Diffuse_ML_interior)), int(nkml), int(nk_rho_varies),
                                  float(ML_KhTr_scale), C.byref(dom), None, None)
    if rc != 0:
        raise RuntimeError("oracle tracer_hordiff failed: %d" % rc)


def _offline_call(fn, h_pre, uhtr, vhtr, dt_offline, num_off_iter, G, GV, CS, Reg, x_before_y, off_ale_mod, min_residual,
                  H_to_kg_m2, L_to_m, kg_m2_to_H, m_to_L, OBC, nthreads):
    g, cs, dom = make_grid(G, GV), make_cs(CS), make_domain(G.Domain)
    tr = make_tracers(Reg)
    obc, _keep = make_obc(OBC)
    conv, nit = C.c_int(0), C.c_int(0)
    res = np.full(num_off_iter + 1, np.nan)
    rc = fn(C.byref(g), C.byref(cs), tr, Reg.ntr, _p(h_pre), _p(uhtr), _p(vhtr), float(dt_offline), int(num_off_iter),
            int(off_ale_mod), int(bool(x_before_y)), float(min_residual), float(H_to_kg_m2), float(L_to_m), float(kg_m2_to_H),
            float(m_to_L), C.byref(dom), C.byref(obc) if obc is not None else None, C.byref(conv), C.byref(nit), _p(res),
            int(nthreads))
    if rc != 0:
        raise RuntimeError("offline advection loop failed: %d" % rc)
    return bool(conv.value), nit.value, res[:nit.value + 1].copy()


def offline_advection_loop(h_pre, uhtr, vhtr, dt_offline, num_off_iter, G, GV, CS, Reg, x_before_y=True, off_ale_mod=10 ** 6,
                           min_residual=0.0, H_to_kg_m2=1.0, L_to_m=1.0, kg_m2_to_H=1.0, m_to_L=1.0, OBC=None, nthreads=1):
    """The iteration loop of offline_advection_ale (MOM_offline_main.F90:254-312, 320-411, 414-415) without its ALE step;
    h_pre, uhtr, vhtr and Reg.Tr[m].t are updated in place.  Returns (converged, iterations run, remaining_transport_sum
    at :312 and after every iteration)."""
    return _offline_call(lib().ora_offline_advection_loop, h_pre, uhtr, vhtr, dt_offline, num_off_iter, G, GV, CS, Reg,
                         x_before_y, off_ale_mod, min_residual, H_to_kg_m2, L_to_m, kg_m2_to_H, m_to_L, OBC, nthreads)


def remaining_transport_sum(uhtr, vhtr, h_new, G, GV, H_to_kg_m2=1.0, L_to_m=1.0, kg_m2_to_H=1.0, m_to_L=1.0) -> float:
    g = make_grid(G, GV)
    return float(lib().ora_remaining_transport_sum(C.byref(g), _p(uhtr), _p(vhtr), _p(h_new), float(H_to_kg_m2), float(L_to_m),
                                                   float(kg_m2_to_H), float(m_to_L)))


def advect_tracer_multi(ranks: Sequence[dict], dom, CS, dt, x_first_in=None, max_iter_in=None,
                        update_vol_prev=None, impl=None) -> List[int]:
    """Emulated multi-PE run.  Each entry of ``ranks`` is a dict with keys
    G, GV, Reg, h_end, uhtr, vhtr and optional vol_prev, uhr_out, vhr_out, OBC (the rank's own ocean_OBC_type)."""
    n = len(ranks)
    arr = (ora_rank * n)()
    keep = []
    for r, d in enumerate(ranks):
        G = d["G"]
        arr[r].G = make_grid(G, d["GV"])
        arr[r].ioff, arr[r].joff = G.ioff, G.joff
        tr = make_tracers(d["Reg"])
        keep.append(tr)
        arr[r].Tr = C.cast(tr, C.POINTER(ora_tracer))
        arr[r].h_end, arr[r].uhtr, arr[r].vhtr = _p(d["h_end"]), _p(d["uhtr"]), _p(d["vhtr"])
        arr[r].vol_prev = _p(d.get("vol_prev"))
        arr[r].uhr_out, arr[r].vhr_out = _p(d.get("uhr_out")), _p(d.get("vhr_out"))
    cs, od = make_cs(CS), make_domain(dom)
    if impl is not None or any(d.get("OBC") is not None for d in ranks):
        obcs = (C.POINTER(ora_obc) * n)()
        for r, d in enumerate(ranks):
            o, k = make_obc(d.get("OBC"))
            keep.append((o, k))
            obcs[r] = C.pointer(o) if o is not None else None
        if impl is not None:      # per-rank routine supplied by the caller (the translated reference, oracle/f2c_ref.py)
            f = lib().ora_advect_tracer_multi_impl
            f.restype = C.c_int
            rc = f(n, arr, C.byref(od), C.byref(cs), ranks[0]["Reg"].ntr, C.c_double(float(dt)), _opt_flag(x_first_in),
                   -1 if max_iter_in is None else int(max_iter_in), _opt_flag(update_vol_prev), obcs,
                   C.cast(impl, C.c_void_p))
            if rc != 0:
                raise RuntimeError("oracle advect_tracer_multi_impl failed: %d" % rc)
            return [arr[r].n_iter for r in range(n)]
        f = lib().ora_advect_tracer_multi_obc
        f.restype = C.c_int
        rc = f(n, arr, C.byref(od), C.byref(cs), ranks[0]["Reg"].ntr, C.c_double(float(dt)), _opt_flag(x_first_in),
               -1 if max_iter_in is None else int(max_iter_in), _opt_flag(update_vol_prev), obcs)
        if rc != 0:
            raise RuntimeError("oracle advect_tracer_multi_obc failed: %d" % rc)
        return [arr[r].n_iter for r in range(n)]
    rc = lib().ora_advect_tracer_multi(n, arr, C.byref(od), C.byref(cs), ranks[0]["Reg"].ntr, float(dt),
                                       _opt_flag(x_first_in), -1 if max_iter_in is None else int(max_iter_in),
                                       _opt_flag(update_vol_prev))
    if rc != 0:
        raise RuntimeError("oracle advect_tracer_multi failed: %d" % rc)
    return [arr[r].n_iter for r in range(n)]


def bitcount_chksum(a: np.ndarray) -> int:
    a = np.ascontiguousarray(a, dtype=np.float64)
    return int(lib().ora_bitcount_chksum(_p(a), a.size))


def efp_sum(a: np.ndarray):
    """Order-invariant extended-fixed-point sum (MOM_coms.F90:508-590).
    Returns (limbs tuple, real value)."""
    a = np.ascontiguousarray(a, dtype=np.float64)
    limbs = (C.c_longlong * 6)()
    lib().ora_efp_sum(_p(a), a.size, limbs)
    return tuple(limbs), float(lib().ora_efp_to_real(limbs))


# ---------------------------------------------------------------------------
# Either side of advect_tracer, restated with numpy.  Arrays are [k, j, i] views of the
# Fortran arrays (i fastest), indices are the grid's 1-based values mapped through G.pi / G.pj / G.pI / G.pJ.  Every
# sum runs over k in the reference's order, element-wise over (i, j): the same IEEE operations as the Fortran loops.
# ---------------------------------------------------------------------------
def accumulate_transports(uhtr, vhtr, uh, vh, dt, G):
    """uhtr(I,j,k) = uhtr(I,j,k) + uh(I,j,k)*dt  (j=js-2,je+2 ; I=Isq-2,Ieq+2) and the same for vhtr
    (/root/reference/src/core/MOM_dynamics_split_RK2.F90:1045-1052; Isq = G%IscB, :409), clipped to the declared bounds."""
    IscB = G.isc - 1 if G.symmetric else G.isc
    JscB = G.jsc - 1 if G.symmetric else G.jsc
    I0, I1 = max(IscB - 2, G.IsdB), min(G.iec + 2, G.ied)
    j0, j1 = max(G.jsc - 2, G.jsd), min(G.jec + 2, G.jed)
    su = (slice(None), slice(G.pj(j0), G.pj(j1) + 1), slice(G.pI(I0), G.pI(I1) + 1))
    uhtr[su] = uhtr[su] + uh[su] * dt
    i0, i1 = max(G.isc - 2, G.isd), min(G.iec + 2, G.ied)
    J0, J1 = max(JscB - 2, G.JsdB), min(G.jec + 2, G.jed)
    sv = (slice(None), slice(G.pJ(J0), G.pJ(J1) + 1), slice(G.pi(i0), G.pi(i1) + 1))
    vhtr[sv] = vhtr[sv] + vh[sv] * dt


def column_diag(kind, G, nz, t=None, h_diag=None, H_to_m=1.0, advection_xy=None):
    """work2d of post_tracer_transport_diagnostics (/root/reference/src/tracer/MOM_tracer_registry.F90:725-815):
    kind 0 adv_xy_2d (:786-791), 1 zint (:796-802), 2 zint_100m (:745-766,805-811).  Returns (jsd:jed, isd:ied)."""
    work = np.zeros((G.jed - G.jsd + 1, G.ied - G.isd + 1))
    sl = (slice(G.pj(G.jsc), G.pj(G.jec) + 1), slice(G.pi(G.isc), G.pi(G.iec) + 1))
    if kind == 0:
        for k in range(nz):
            work[sl] = work[sl] + advection_xy[k][sl]
    elif kind == 1:
        for k in range(nz):
            work[sl] = work[sl] + (h_diag[k][sl] * H_to_m) * t[k][sl]
    else:
        frac = np.zeros((nz,) + work[sl].shape)
        zbot = np.zeros(work[sl].shape)
        khi = 0
        for k in range(nz):
            ztop = zbot
            zbot = ztop + h_diag[k][sl] * H_to_m
            with np.errstate(divide="ignore", invalid="ignore"):
                part = (100.0 - ztop) / (zbot - ztop)
            frac[k] = np.where(zbot <= 100.0, 1.0, np.where(ztop < 100.0, part, 0.0))
            if np.any(frac[k] > 0):
                khi = k + 1
        for k in range(khi):
            work[sl] = work[sl] + frac[k] * ((h_diag[k][sl] * H_to_m) * t[k][sl])
    return work
