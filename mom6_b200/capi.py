"""ctypes binding of the C ABI in include/mom6adv.h (libmom6adv_b200.so).

This is exactly what the Fortran shim (fortran/MOM_tracer_advect.F90) binds
through ISO_C_BINDING; the Python host mirror goes through the same entry
points, so tests that call this module exercise the drop-in boundary itself.
There is no CPU fallback: a missing library raises at load time.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# MOM6ADV_LIB selects a differently-flagged build of the same sources (kernel A/B timing, scripts/variants.py)
LIB_PATH = os.environ.get("MOM6ADV_LIB") or os.path.join(_HERE, "libmom6adv_b200.so")

OK, ERR_NOT_INIT, ERR_INVALID, ERR_UNSUPPORTED, ERR_CUDA, ERR_COMM = range(6)
F_H_END, F_UHTR, F_VHTR, F_VOL_PREV, F_UHR, F_VHR, F_HPREV = range(7)
CD_ADV_XY_2D, CD_ZINT, CD_ZINT_100M = range(3)
KC_NAMES = ("prologue", "halo", "flags", "mass", "tracer_x", "tracer_y", "other", "sparse",
            "exch_pack", "exch_wire", "exch_unpack")
DIAG_KINDS = ("ad_x", "ad_y", "ad2d_x", "ad2d_y", "advection_xy")

_dp = C.POINTER(C.c_double)


class mom6adv_grid(C.Structure):
    _fields_ = [(n, C.c_int) for n in
                ("isc", "iec", "jsc", "jec", "isd", "ied", "jsd", "jed",
                 "IsdB", "IedB", "JsdB", "JedB", "nk")] + \
               [(n, _dp) for n in ("areaT", "IareaT", "mask2dT", "mask2dCu", "mask2dCv")] + \
               [("Angstrom_H", C.c_double), ("H_subroundoff", C.c_double)]


class mom6adv_config(C.Structure):
    _fields_ = [("dt_baroclinic", C.c_double)] + \
               [(n, C.c_int) for n in ("usePPM", "useHuynh", "useHuynhStencilBug", "debug", "first_direction")]


class mom6adv_domain(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("nig", "njg", "npi", "npj", "pi", "pj", "ioff", "joff",
                                       "cyclic_x", "cyclic_y", "tripolar_n")]


class mom6adv_obc_tracer(C.Structure):
    _fields_ = [("ntr_index", C.c_int), ("tres", _dp), ("OBC_inflow_conc", C.c_double)]


class mom6adv_obc_segment(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("is_E_or_W", "is_N_or_S", "direction", "specified",
                                       "isd", "ied", "jsd", "jed", "IsdB", "IedB", "JsdB", "JedB", "ntseg")] + \
               [("Tr", C.POINTER(mom6adv_obc_tracer))]


class mom6adv_obc(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("OBC_pe", "number_of_segments", "open_u_BCs_exist_globally",
                                       "open_v_BCs_exist_globally", "specified_u_BCs_exist_globally",
                                       "specified_v_BCs_exist_globally")] + \
               [("segment", C.POINTER(mom6adv_obc_segment))]


def make_obc(OBC):
    """ocean_OBC_type (mom6_b200.types.OceanOBC) -> mom6adv_obc; returns (struct, keep-alive list)."""
    keep = []
    segs = (mom6adv_obc_segment * max(OBC.number_of_segments, 1))()
    for n, sg in enumerate(OBC.segment):
        s = segs[n]
        s.is_E_or_W, s.is_N_or_S = int(sg.is_E_or_W), int(sg.is_N_or_S)
        s.direction, s.specified = int(sg.direction), int(sg.specified)
        for f in ("isd", "ied", "jsd", "jed", "IsdB", "IedB", "JsdB", "JedB"):
            setattr(s, f, int(getattr(sg.HI, f)))
        if sg.tr_Reg is None:
            s.ntseg = -1
        else:
            s.ntseg = len(sg.tr_Reg)
            trs = (mom6adv_obc_tracer * max(len(sg.tr_Reg), 1))()
            for m, t in enumerate(sg.tr_Reg):
                trs[m].ntr_index, trs[m].OBC_inflow_conc = int(t.ntr_index), float(t.OBC_inflow_conc)
                if t.tres is not None:
                    assert t.tres.dtype == np.float64 and t.tres.flags["C_CONTIGUOUS"]
                    trs[m].tres = t.tres.ctypes.data_as(_dp)
            keep.append(trs)
            s.Tr = C.cast(trs, C.POINTER(mom6adv_obc_tracer))
    o = mom6adv_obc()
    o.OBC_pe, o.number_of_segments = int(OBC.OBC_pe), OBC.number_of_segments
    o.open_u_BCs_exist_globally, o.open_v_BCs_exist_globally = int(OBC.open_u_BCs_exist_globally), int(OBC.open_v_BCs_exist_globally)
    o.specified_u_BCs_exist_globally = int(OBC.specified_u_BCs_exist_globally)
    o.specified_v_BCs_exist_globally = int(OBC.specified_v_BCs_exist_globally)
    o.segment = C.cast(segs, C.POINTER(mom6adv_obc_segment))
    keep.append(segs)
    return o, keep


# every symbol include/mom6adv.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "mom6adv_plan_create": (C.c_int, [C.POINTER(mom6adv_grid), C.POINTER(mom6adv_config),
                                      C.POINTER(mom6adv_domain), C.c_int, C.POINTER(_P)]),
    "mom6adv_plan_destroy": (C.c_int, [_P]),
    "mom6adv_last_error": (C.c_char_p, [_P]),
    "mom6adv_version": (C.c_char_p, []),
    "mom6adv_comm_unique_id": (C.c_int, [C.c_char_p]),
    "mom6adv_comm_init": (C.c_int, [_P, C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                    C.POINTER(C.c_int)]),
    "mom6adv_exchange_list": (C.c_int, [C.POINTER(mom6adv_grid), C.POINTER(mom6adv_domain), C.POINTER(C.c_int),
                                        C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int, C.c_int,
                                        C.c_int, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_int)]),
    "mom6adv_hordiff": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int]),
    "mom6adv_offline_advect": (C.c_int, [_P, C.c_double, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                                         C.c_double, C.c_double, C.POINTER(C.c_int), C.POINTER(C.c_int), _P]),
    "mom6adv_bind_tracer_df": (C.c_int, [_P, C.c_int, _P, _P, _P, _P]),
    "mom6adv_hordiff_diag": (C.c_int, [_P, _P, _P, _P, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double]),
    "mom6adv_set_obc": (C.c_int, [_P, C.POINTER(mom6adv_obc)]),
    "mom6adv_set_ntr": (C.c_int, [_P, C.c_int]),
    "mom6adv_bind_tracer": (C.c_int, [_P, C.c_int, _P, C.c_double, _P, _P, _P, _P, _P]),
    "mom6adv_advect": (C.c_int, [_P, _P, _P, _P, C.c_double, C.c_int, _P, C.c_int, C.c_int, _P, _P,
                                 C.POINTER(C.c_int)]),
    "mom6adv_upload_tracer": (C.c_int, [_P, C.c_int, _P]),
    "mom6adv_download_tracer": (C.c_int, [_P, C.c_int, _P]),
    "mom6adv_upload_field": (C.c_int, [_P, C.c_int, _P]),
    "mom6adv_download_field": (C.c_int, [_P, C.c_int, _P]),
    "mom6adv_download_diag": (C.c_int, [_P, C.c_int, C.c_int, _P]),
    "mom6adv_advect_resident": (C.c_int, [_P, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.POINTER(C.c_int)]),
    "mom6adv_sync": (C.c_int, [_P]),
    "mom6adv_accumulate_transports": (C.c_int, [_P, _P, _P, C.c_double]),
    "mom6adv_zero_transports": (C.c_int, [_P]),
    "mom6adv_column_diag": (C.c_int, [_P, C.c_int, C.c_int, _P, C.c_double, _P]),
    "mom6adv_zint100_noalloc": (C.c_int, [_P, _P, C.c_int, C.c_int, C.c_int, C.c_double, _P, _P, C.c_char_p, C.c_int]),
    "mom6adv_last_advect_ms": (C.c_double, [_P]),
    "mom6adv_kernel_launches": (C.c_longlong, [_P]),
    "mom6adv_device_bytes": (C.c_longlong, [_P]),
    "mom6adv_halo_bytes": (C.c_longlong, [_P]),
    "mom6adv_set_option": (C.c_int, [_P, C.c_char_p, C.c_int]),
    "mom6adv_kernel_class_time": (C.c_int, [_P, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_longlong)]),
    "mom6adv_stream": (C.c_void_p, [_P]),
}

_lib = None


def lib():
    """Load libmom6adv_b200.so (built in-tree by mom6_b200.build).  Raises if missing."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("%s is missing: run `python -m mom6_b200.build` (there is no CPU fallback)" % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(l, name)     # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        _lib = l
    return _lib


class Mom6AdvError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__("mom6adv status %d: %s" % (status, msg))
        self.status = status


def _order_after_torch(*arrays) -> None:
    """The plan works on its own non-blocking stream.  A torch CUDA tensor handed to it may still be being written by
    kernels queued on torch's stream (and, once the caller drops it, its memory may be handed to the next tensor):
    wait for torch's stream before the library reads or writes such a tensor."""
    for a in arrays:
        if a is not None and getattr(a, "is_cuda", False):
            import torch
            torch.cuda.current_stream(a.device).synchronize()
            return


def ptr(a) -> Optional[int]:
    """Raw address of a numpy array / torch tensor / int / None."""
    if a is None:
        return None
    if isinstance(a, int):
        return a
    if isinstance(a, np.ndarray):
        if not (a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]):
            raise ValueError("need a contiguous float64 array")
        return a.ctypes.data
    if hasattr(a, "data_ptr"):   # torch tensor (host or device)
        if not a.is_contiguous() or a.element_size() != 8:
            raise ValueError("need a contiguous float64 tensor")
        return a.data_ptr()
    raise TypeError("unsupported array type %r" % type(a))


def _iarr(v):
    return (C.c_int * len(v))(*[int(x) for x in v])


def make_structs(G, GV, CS=None):
    """mom6adv_grid / mom6adv_config / mom6adv_domain PODs of a host grid (plus the arrays to keep alive)."""
    keep = []
    g = mom6adv_grid()
    for n in ("isc", "iec", "jsc", "jec", "isd", "ied", "jsd", "jed", "IsdB", "IedB", "JsdB", "JedB"):
        setattr(g, n, int(getattr(G, n)))
    g.nk = int(GV.ke)
    for n in ("areaT", "IareaT", "mask2dT", "mask2dCu", "mask2dCv"):
        a = np.ascontiguousarray(getattr(G, n), dtype=np.float64)
        keep.append(a)
        setattr(g, n, a.ctypes.data_as(_dp))
    g.Angstrom_H, g.H_subroundoff = GV.Angstrom_H, GV.H_subroundoff
    c = mom6adv_config()
    if CS is not None:
        c.dt_baroclinic = CS.dt
        c.usePPM, c.useHuynh, c.useHuynhStencilBug = int(CS.usePPM), int(CS.useHuynh), int(CS.useHuynhStencilBug)
        c.debug = int(getattr(CS, "debug", False))
    c.first_direction = int(G.first_direction)
    D = G.Domain
    d = mom6adv_domain()
    d.nig, d.njg = D.nig, D.njg
    d.npi, d.npj = D.layout
    d.pi, d.pj = D.pe
    d.ioff, d.joff = G.ioff, G.joff
    d.cyclic_x, d.cyclic_y, d.tripolar_n = int(D.cyclic_x), int(D.cyclic_y), int(D.tripolar_n)
    return g, c, d, keep


def comm_unique_id() -> bytes:
    buf = C.create_string_buffer(128)
    rc = lib().mom6adv_comm_unique_id(buf)
    if rc != OK:
        raise Mom6AdvError(rc, lib().mom6adv_last_error(None).decode())
    return buf.raw


def exchange_list(G, GV, extents, peer: int, pos: int, what: int):
    """Host-only: one list of the exchange plan of G's rank (see mom6adv_exchange_list).  Returns (list, geom)."""
    g, c, d, keep = make_structs(G, GV)
    ib, ie, jb, je = (_iarr(v) for v in extents)
    geom = (C.c_int * 4)()
    n = lib().mom6adv_exchange_list(C.byref(g), C.byref(d), ib, ie, jb, je, peer, pos, what, None, 0, geom)
    if n < 0:
        raise Mom6AdvError(ERR_INVALID, "exchange_list failed")
    out = (C.c_int * max(n, 1))()
    lib().mom6adv_exchange_list(C.byref(g), C.byref(d), ib, ie, jb, je, peer, pos, what, out, n, geom)
    return np.array(out[:n], dtype=np.int64), tuple(geom)


class Plan:
    """Thin object wrapper around mom6adv_plan*."""

    def __init__(self, G, GV, CS, device: int = -1):
        self._l = lib()
        self._h = _P()
        self._keep = []
        g, c, d, keep = make_structs(G, GV, CS)
        self._keep += keep
        rc = self._l.mom6adv_plan_create(C.byref(g), C.byref(c), C.byref(d), int(device), C.byref(self._h))
        if rc != OK:
            raise Mom6AdvError(rc, self._l.mom6adv_last_error(None).decode())
        self.G, self.GV, self.ntr = G, GV, 0

    # -- helpers ---------------------------------------------------------------
    def _ck(self, rc):
        if rc != OK:
            raise Mom6AdvError(rc, self._l.mom6adv_last_error(self._h).decode())

    def destroy(self):
        if self._h:
            self._l.mom6adv_plan_destroy(self._h)
            self._h = _P()

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass

    # -- multi-GPU --------------------------------------------------------------
    def comm_init(self, id128: Optional[bytes], extents):
        """extents = (ibeg[npi], iend[npi], jbeg[npj], jend[npj]) global compute extents of the layout."""
        ib, ie, jb, je = (_iarr(v) for v in extents)
        self._ck(self._l.mom6adv_comm_init(self._h, id128, ib, ie, jb, je))

    # -- registry ---------------------------------------------------------------
    def set_ntr(self, ntr: int):
        self._ck(self._l.mom6adv_set_ntr(self._h, ntr))
        self.ntr = ntr

    def set_obc(self, OBC):
        """The OBC argument of advect_tracer: None == unassociated."""
        if OBC is None:
            self._ck(self._l.mom6adv_set_obc(self._h, None))
            return
        o, keep = make_obc(OBC)
        self._ck(self._l.mom6adv_set_obc(self._h, C.byref(o)))

    def bind_tracer(self, m, t, conc_underflow=0.0, ad_x=None, ad_y=None, ad2d_x=None, ad2d_y=None,
                    advection_xy=None):
        self._ck(self._l.mom6adv_bind_tracer(self._h, m, ptr(t), float(conc_underflow), ptr(ad_x), ptr(ad_y),
                                             ptr(ad2d_x), ptr(ad2d_y), ptr(advection_xy)))

    # -- MIRROR ----------------------------------------------------------------
    def advect(self, h_end, uhtr, vhtr, dt, x_first=-1, vol_prev=None, max_iter_in=-1, update_vol_prev=-1,
               uhr_out=None, vhr_out=None) -> int:
        n = C.c_int(0)
        _order_after_torch(h_end, uhtr, vhtr, vol_prev, uhr_out, vhr_out)
        self._ck(self._l.mom6adv_advect(self._h, ptr(h_end), ptr(uhtr), ptr(vhtr), float(dt), int(x_first),
                                        ptr(vol_prev), int(max_iter_in), int(update_vol_prev), ptr(uhr_out),
                                        ptr(vhr_out), C.byref(n)))
        return n.value

    # -- RESIDENT --------------------------------------------------------------
    # (each of these returns after the plan's stream has finished with the array)
    def upload_tracer(self, m, src):
        _order_after_torch(src); self._ck(self._l.mom6adv_upload_tracer(self._h, m, ptr(src)))
    def download_tracer(self, m, dst):
        _order_after_torch(dst); self._ck(self._l.mom6adv_download_tracer(self._h, m, ptr(dst)))
    def upload_field(self, f, src):
        _order_after_torch(src); self._ck(self._l.mom6adv_upload_field(self._h, f, ptr(src)))
    def download_field(self, f, dst):
        _order_after_torch(dst); self._ck(self._l.mom6adv_download_field(self._h, f, ptr(dst)))
    def download_diag(self, m, kind, dst):
        _order_after_torch(dst); self._ck(self._l.mom6adv_download_diag(self._h, m, kind, ptr(dst)))

    def advect_resident(self, dt, x_first=-1, max_iter_in=-1, use_vol_prev=0, update_vol_prev=0) -> int:
        n = C.c_int(0)
        self._ck(self._l.mom6adv_advect_resident(self._h, float(dt), int(x_first), int(max_iter_in),
                                                 int(use_vol_prev), int(update_vol_prev), C.byref(n)))
        return n.value

    def sync(self): self._ck(self._l.mom6adv_sync(self._h))

    # -- either side of advect_tracer --------------------------
    def accumulate_transports(self, uh, vh, dt):
        """uhtr += uh*dt ; vhtr += vh*dt on the resident fields (MOM_dynamics_split_RK2.F90:1045-1052)."""
        _order_after_torch(uh, vh)
        self._ck(self._l.mom6adv_accumulate_transports(self._h, ptr(uh), ptr(vh), float(dt)))

    def hordiff(self, khdt_x, khdt_y, num_itts, h_neglect, h=None, Diffuse_ML_interior=False, nkml=0, nk_rho_varies=0,
                ML_KhTr_scale=0.0, want_df=False):
        """The diffusion iteration of tracer_hordiff (MOM_tracer_hor_diff.F90:539-607) on the resident tracers; h=None
        uses the resident H_END field (MOM.F90:1441 passes advect_tracer's h_end)."""
        _order_after_torch(khdt_x, khdt_y, h)
        self._ck(self._l.mom6adv_hordiff(self._h, ptr(h) if h is not None else None, ptr(khdt_x), ptr(khdt_y),
                                         int(num_itts), float(h_neglect), int(bool(Diffuse_ML_interior)), int(nkml),
                                         int(nk_rho_varies), float(ML_KhTr_scale), int(bool(want_df))))

    def bind_tracer_df(self, m, df_x=None, df_y=None, df2d_x=None, df2d_y=None):
        """Which of tracer_hordiff's diffusive-flux diagnostics tracer m has (non-None = associated; values stay in HBM,
        download_diag kinds 5..8)."""
        self._ck(self._l.mom6adv_bind_tracer_df(self._h, int(m), ptr(df_x), ptr(df_y), ptr(df2d_x), ptr(df2d_y)))

    def hordiff_diag(self, khdt_x, khdt_y, num_itts, h_neglect, dt, h=None, Diffuse_ML_interior=False, nkml=0,
                     nk_rho_varies=0, ML_KhTr_scale=0.0):
        """hordiff() with the df_* diagnostics of the tracers bound by bind_tracer_df (MOM_tracer_hor_diff.F90:576-591).
        Needs the option hordiff_df: the path has not run on a GPU yet."""
        _order_after_torch(khdt_x, khdt_y, h)
        self._ck(self._l.mom6adv_hordiff_diag(self._h, ptr(h) if h is not None else None, ptr(khdt_x), ptr(khdt_y),
                                              int(num_itts), float(h_neglect), int(bool(Diffuse_ML_interior)), int(nkml),
                                              int(nk_rho_varies), float(ML_KhTr_scale), float(dt)))

    def offline_advect(self, dt_offline, num_off_iter, x_before_y=True, off_ale_mod=10 ** 6, min_residual=0.0,
                       H_to_kg_m2=1.0, L_to_m=1.0, kg_m2_to_H=1.0, m_to_L=1.0):
        """The iteration loop of offline_advection_ale (MOM_offline_main.F90:254-415) on the resident fields.
        Returns (converged, trips, remaining_transport_sum before the loop and after every trip)."""
        conv, nit = C.c_int(0), C.c_int(0)
        res = np.full(int(num_off_iter) + 1, np.nan)
        self._ck(self._l.mom6adv_offline_advect(self._h, float(dt_offline), int(num_off_iter), int(off_ale_mod),
                                                int(bool(x_before_y)), float(min_residual), float(H_to_kg_m2), float(L_to_m),
                                                float(kg_m2_to_H), float(m_to_L), C.byref(conv), C.byref(nit),
                                                res.ctypes.data))
        return bool(conv.value), nit.value, res[:nit.value + 1].copy()

    def zero_transports(self):
        """CS%uhtr(:,:,:) = 0.0 ; CS%vhtr(:,:,:) = 0.0 (MOM.F90:1465-1466)."""
        self._ck(self._l.mom6adv_zero_transports(self._h))

    def column_diag(self, m, kind, h_diag, H_to_m, out2d):
        """Vertical sums of post_tracer_transport_diagnostics (MOM_tracer_registry.F90:725-815) for resident tracer m."""
        _order_after_torch(h_diag, out2d)
        self._ck(self._l.mom6adv_column_diag(self._h, int(m), int(kind), ptr(h_diag), float(H_to_m), ptr(out2d)))
    def set_option(self, key: str, value: int): self._ck(self._l.mom6adv_set_option(self._h, key.encode(), value))

    def kernel_class_time(self, kclass: int):
        ms, n = C.c_double(0.0), C.c_longlong(0)
        self._ck(self._l.mom6adv_kernel_class_time(self._h, kclass, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    @property
    def stream(self) -> int: return int(self._l.mom6adv_stream(self._h) or 0)

    @property
    def last_advect_ms(self) -> float: return float(self._l.mom6adv_last_advect_ms(self._h))
    @property
    def kernel_launches(self) -> int: return int(self._l.mom6adv_kernel_launches(self._h))
    @property
    def device_bytes(self) -> int: return int(self._l.mom6adv_device_bytes(self._h))
    @property
    def halo_bytes(self) -> int: return int(self._l.mom6adv_halo_bytes(self._h))


NOALLOC_MAX_COLUMNS = 8388608


def zint100_noalloc(h_diag, t, ni: int, nj: int, nk: int, H_to_m: float, out2d, stream: int = 0) -> None:
    """mom6adv_zint100_noalloc: the 100 m column integral on the caller's DEVICE arrays (ni, nj, nk) -> (ni, nj), with
    the library's fixed scratch and no allocation (no counterpart in the reference; see include/mom6adv.h)."""
    if not stream:
        pass      # default stream: ordered with torch's default stream by CUDA itself
    else:
        _order_after_torch(h_diag, t, out2d)
    buf = C.create_string_buffer(256)
    rc = lib().mom6adv_zint100_noalloc(ptr(h_diag), ptr(t), int(ni), int(nj), int(nk), float(H_to_m), ptr(out2d),
                                       stream or None, buf, 256)
    if rc != 0:
        raise Mom6AdvError(rc, buf.value.decode() or "mom6adv_zint100_noalloc failed")
