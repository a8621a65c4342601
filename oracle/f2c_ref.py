"""TEST INFRASTRUCTURE ONLY -- ctypes wrapper around oracle/_ref/libref_f2c*.so: the UNMODIFIED reference file
/root/reference/src/tracer/MOM_tracer_advect.F90 translated statement by statement to C++ by oracle/f2c/f2cpp.py
(there is no Fortran compiler in the image) and compiled with g++ -O2 -ffp-contract=off.

build() runs the translator only where /root/reference exists (the build container); the GPU box uses the prebuilt
libraries that travel with the snapshot (oracle/_ref/ is git-ignored, not gpurun-ignored).  Nothing of the reference
is stored in the repository.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

from . import oracle as ora

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
_SRC = "/root/reference/src/tracer/MOM_tracer_advect.F90"
_libs = {}

VARIANTS = {"": "libref_f2c.so", "lt": "libref_f2c_lt.so", "mm": "libref_f2c_mm.so", "chk": "libref_f2c_chk.so"}
_EXTRA = ("libref_f2c_hordiff.so", "libref_f2c_offline.so", "libref_f2c_step.so")


def available() -> bool:
    return os.path.exists(os.path.join(_REF_DIR, VARIANTS[""])) or os.path.exists(_SRC)


def build(force: bool = False) -> bool:
    """Translate + compile when the reference source is present; returns whether the libraries exist."""
    ora.build()
    have = all(os.path.exists(os.path.join(_REF_DIR, v)) for v in list(VARIANTS.values()) + list(_EXTRA))
    if os.path.exists(_SRC) and (force or not have):
        if force:
            subprocess.run(["make", "-C", os.path.join(_HERE, "f2c"), "clean"], check=True, capture_output=True)
        subprocess.run(["make", "-C", os.path.join(_HERE, "f2c")], check=True, capture_output=True)
        have = True
    elif os.path.exists(_SRC):
        subprocess.run(["make", "-C", os.path.join(_HERE, "f2c")], check=True, capture_output=True)     # up to date?
    return have


def lib(variant: str = ""):
    if variant not in _libs:
        if not build():
            raise RuntimeError("oracle/_ref/libref_f2c*.so missing and /root/reference absent: cannot build the translated reference")
        ora.lib()      # liboracle_advect.so (group-pass emulation) first: the rpath of the ref library points at it
        l = C.CDLL(os.path.join(_REF_DIR, VARIANTS[variant]))
        l.ref_advect_tracer_obc.restype = C.c_int
        l.ref_advect_tracer_obc.argtypes = [
            C.POINTER(ora.ora_grid), C.POINTER(ora.ora_cs), C.POINTER(ora.ora_tracer), C.c_int,
            ora._dp, ora._dp, ora._dp, C.c_double, C.c_int, ora._dp, C.c_int, C.c_int, ora._dp, ora._dp,
            C.POINTER(ora.ora_domain), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int), C.POINTER(ora.ora_obc)]
        _libs[variant] = l
    return _libs[variant]


def advect_tracer(h_end, uhtr, vhtr, OBC, dt, G, GV, US, CS, Reg, x_first_in=None, vol_prev=None, max_iter_in=None,
                  update_vol_prev=None, uhr_out=None, vhr_out=None, variant: str = "", pass_fn=None, sum_fn=None) -> int:
    """The translated reference's advect_tracer (same argument list as oracle.advect_tracer).  Returns the number of
    group passes it made (the reference does not report its iteration count)."""
    g, cs, dom = ora.make_grid(G, GV), ora.make_cs(CS), ora.make_domain(G.Domain)
    tr = ora.make_tracers(Reg)
    npass = C.c_int(0)
    pf = C.cast(pass_fn, C.c_void_p) if pass_fn is not None else None
    sf = C.cast(sum_fn, C.c_void_p) if sum_fn is not None else None
    obc, _keep = ora.make_obc(OBC)
    rc = lib(variant).ref_advect_tracer_obc(
        C.byref(g), C.byref(cs), tr, Reg.ntr, ora._p(h_end), ora._p(uhtr), ora._p(vhtr), float(dt),
        ora._opt_flag(x_first_in), ora._p(vol_prev), -1 if max_iter_in is None else int(max_iter_in),
        ora._opt_flag(update_vol_prev), ora._p(uhr_out), ora._p(vhr_out), C.byref(dom), pf, sf, None,
        C.byref(npass), C.byref(obc) if obc is not None else None)
    if rc != 0:
        raise RuntimeError("translated reference advect_tracer failed (FATAL): %d" % rc)
    return npass.value


_hd = None


def tracer_hordiff(h, khdt_x, khdt_y, num_itts, dt, G, GV, Reg, Diffuse_ML_interior=False, nkml=0, nk_rho_varies=0,
                   ML_KhTr_scale=0.0, df=None) -> None:
    """The translated lines MOM_tracer_hor_diff.F90:539-607 (oracle/f2c/hordiff_wrap.cpp, bounds-checked build)."""
    global _hd
    if _hd is None:
        lib()      # builds everything and loads libref_f2c.so, which the hordiff library links against
        _hd = C.CDLL(os.path.join(_REF_DIR, "libref_f2c_hordiff.so"))
        _hd.ref_tracer_hordiff.restype = C.c_int
        _hd.ref_tracer_hordiff.argtypes = [C.POINTER(ora.ora_grid), C.POINTER(ora.ora_tracer), C.c_int, ora._dp, ora._dp,
                                           ora._dp, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_double,
                                           C.c_double, C.POINTER(ora.ora_domain), C.c_void_p, C.c_void_p]
    g, dom = ora.make_grid(G, GV), ora.make_domain(G.Domain)
    tr = ora.make_tracers(Reg)
    if df is not None:      # the df_x / df_y / df2d_x / df2d_y diagnostics of :576-591 (same dict as oracle.tracer_hordiff)
        tabs = ora.df_pointer_tables(df, Reg.ntr)
        f = _hd.ref_tracer_hordiff_df
        f.restype = C.c_int
        rc = f(C.byref(g), tr, C.c_int(Reg.ntr), ora._p(h), ora._p(khdt_x), ora._p(khdt_y), C.c_int(int(num_itts)),
               C.c_double(float(GV.H_subroundoff)), C.c_int(int(bool(Diffuse_ML_interior))), C.c_int(int(nkml)),
               C.c_int(int(nk_rho_varies)), C.c_double(float(ML_KhTr_scale)), C.c_double(float(dt)), C.byref(dom), None, None,
               *tabs)
        if rc != 0:
            raise RuntimeError("translated tracer_hordiff loop (df) failed: %d" % rc)
        return
    rc = _hd.ref_tracer_hordiff(C.byref(g), tr, Reg.ntr, ora._p(h), ora._p(khdt_x), ora._p(khdt_y), int(num_itts),
                                float(GV.H_subroundoff), int(bool(Diffuse_ML_interior)), int(nkml), int(nk_rho_varies),
                                float(ML_KhTr_scale), float(dt), C.byref(dom), None, None)
    if rc != 0:
        raise RuntimeError("translated tracer_hordiff loop failed: %d" % rc)


_off = None


def offline_advection_loop(h_pre, uhtr, vhtr, dt_offline, num_off_iter, G, GV, CS, Reg, x_before_y=True, off_ale_mod=10 ** 6,
                           min_residual=0.0, H_to_kg_m2=1.0, L_to_m=1.0, kg_m2_to_H=1.0, m_to_L=1.0, OBC=None):
    """The translated lines MOM_offline_main.F90:270-415 and :634-648 around the translated advect_tracer
    (oracle/f2c/offline_wrap.cpp, bounds-checked build).  Same returns as oracle.offline_advection_loop."""
    global _off
    if _off is None:
        lib()
        _off = C.CDLL(os.path.join(_REF_DIR, "libref_f2c_offline.so"))
        _off.ref_offline_advection_loop.restype = C.c_int
        _off.ref_offline_advection_loop.argtypes = ora._offline_argtypes()
    return ora._offline_call(_off.ref_offline_advection_loop, h_pre, uhtr, vhtr, dt_offline, num_off_iter, G, GV, CS, Reg,
                             x_before_y, off_ale_mod, min_residual, H_to_kg_m2, L_to_m, kg_m2_to_H, m_to_L, OBC, 1)


def hordiff_num_itts(khdt_x, khdt_y, G, GV, check_diffusive_CFL=True, max_diff_CFL=-1.0):
    """(num_itts, I_numitts) as the translated lines MOM_tracer_hor_diff.F90:368-387 compute them (one PE)."""
    global _hd
    if _hd is None:
        lib()
        _hd = C.CDLL(os.path.join(_REF_DIR, "libref_f2c_hordiff.so"))
        _hd.ref_tracer_hordiff.restype = C.c_int
        _hd.ref_tracer_hordiff.argtypes = [C.POINTER(ora.ora_grid), C.POINTER(ora.ora_tracer), C.c_int, ora._dp, ora._dp,
                                           ora._dp, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_double,
                                           C.c_double, C.POINTER(ora.ora_domain), C.c_void_p, C.c_void_p]
    g = ora.make_grid(G, GV)
    n, inv = C.c_int(0), C.c_double(0.0)
    f = _hd.ref_hordiff_num_itts
    f.restype = C.c_int
    rc = f(C.byref(g), ora._p(khdt_x), ora._p(khdt_y), C.c_int(int(bool(check_diffusive_CFL))), C.c_double(float(max_diff_CFL)),
           C.byref(n), C.byref(inv))
    if rc != 0:
        raise RuntimeError("translated num_itts lines failed: %d" % rc)
    return n.value, inv.value


_step = None


def _step_lib():
    global _step
    if _step is None:
        lib()
        _step = C.CDLL(os.path.join(_REF_DIR, "libref_f2c_step.so"))
        _step.ref_accumulate_transports.restype = C.c_int
        _step.ref_accumulate_transports.argtypes = [C.POINTER(ora.ora_grid), ora._dp, ora._dp, ora._dp, ora._dp, C.c_double]
        _step.ref_column_diag.restype = C.c_int
        _step.ref_column_diag.argtypes = [C.POINTER(ora.ora_grid), C.c_int, ora._dp, ora._dp, C.c_double, ora._dp,
                                          C.POINTER(C.c_int)]
    return _step


def accumulate_transports(uhtr, vhtr, uh, vh, dt, G, GV) -> None:
    """uhtr += uh*dt ; vhtr += vh*dt by the translated lines MOM_dynamics_split_RK2.F90:1045-1053 (in place)."""
    g = ora.make_grid(G, GV)
    rc = _step_lib().ref_accumulate_transports(C.byref(g), ora._p(uhtr), ora._p(vhtr), ora._p(uh), ora._p(vh), float(dt))
    if rc != 0:
        raise RuntimeError("translated transport accumulation failed: %d" % rc)


def column_diag(kind, G, GV, field3d, h_diag, H_to_m=1.0):
    """work2d of post_tracer_transport_diagnostics by the translated lines MOM_tracer_registry.F90:744-767, 785-788,
    795-798, 804-807: kind 0 adv_xy_2d (field3d = advection_xy), 1 zint, 2 zint_100m (field3d = Tr%t).
    Returns (work2d over (jsd:jed, isd:ied), khi)."""
    import numpy as np
    g = ora.make_grid(G, GV)
    work = np.zeros(G.shape_h())
    khi = C.c_int(0)
    hd = h_diag if h_diag is not None else np.zeros_like(field3d)
    rc = _step_lib().ref_column_diag(C.byref(g), int(kind), ora._p(field3d), ora._p(hd), float(H_to_m), ora._p(work), C.byref(khi))
    if rc != 0:
        raise RuntimeError("translated column diagnostics failed: %d" % rc)
    return work, khi.value


def advect_tracer_multi(ranks, dom, CS, dt, x_first_in=None, max_iter_in=None, update_vol_prev=None, variant: str = ""):
    """The emulated multi-PE run of oracle.advect_tracer_multi with the TRANSLATED REFERENCE as every PE's advect_tracer
    (same rank dicts; "OBC" per rank optional).  Returns the number of group passes each PE made."""
    return ora.advect_tracer_multi(ranks, dom, CS, dt, x_first_in=x_first_in, max_iter_in=max_iter_in,
                                   update_vol_prev=update_vol_prev, impl=lib(variant).ref_advect_tracer_obc)
