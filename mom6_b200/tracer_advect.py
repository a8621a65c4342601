"""Host-side mirror of module MOM_tracer_advect
(/root/reference/src/tracer/MOM_tracer_advect.F90:25-40, 52-53, 1090-1158).

Same public names, argument order and error behaviour as the reference:

    tracer_advect_init(Time, G, US, param_file, diag, CS)   :1090
    advect_tracer(h_end, uhtr, vhtr, OBC, dt, G, GV, US, CS, Reg, x_first_in,
                  vol_prev, max_iter_in, update_vol_prev, uhr_out, vhr_out)   :52
    tracer_advect_end(CS)                                   :1153

Everything below the argument marshalling runs in the CUDA library through the
C ABI of include/mom6adv.h; there is no CPU fallback -- a missing library or a
missing GPU raises.
"""
from __future__ import annotations

import math
import warnings
from dataclasses import dataclass, field
from typing import Any, Optional

import numpy as np


class MOMFatalError(RuntimeError):
    """MOM_error(FATAL, ...)"""


@dataclass
class TracerAdvectCS:
    """tracer_advect_CS, MOM_tracer_advect.F90:30-40."""
    dt: float = 0.0                    # baroclinic dynamics time step [T]
    diag: Any = None
    debug: bool = False
    usePPM: bool = False
    useHuynh: bool = False
    useHuynhStencilBug: bool = False
    # GPU plan cache (stands in for pass_uhr_vhr_t_hprev, :39)
    _plan: Any = field(default=None, repr=False)
    # mom6adv_set_option(key, value) pairs applied when the plan is created (tests / experiments)
    gpu_options: dict = field(default_factory=dict, repr=False)

    @property
    def stencil(self) -> int:
        # :118,129-130
        return 3 if (self.usePPM and not self.useHuynhStencilBug) else 2


@dataclass
class UnitScale:
    """unit_scale_type: only s_to_T is used (:1114)."""
    s_to_T: float = 1.0


def tracer_advect_init(Time, G, US, param_file: dict, diag, CS: Optional[TracerAdvectCS]) -> TracerAdvectCS:
    """tracer_advect_init, MOM_tracer_advect.F90:1090-1150.  ``param_file`` is a
    mapping standing in for MOM_input (get_param keys DT, DEBUG,
    TRACER_ADVECTION_SCHEME, USE_HUYNH_STENCIL_BUG).  Returns the control
    structure (the reference allocates it through the pointer argument)."""
    if CS is not None:   # :1103-1106
        warnings.warn("tracer_advect_init called with associated control structure.")
        return CS
    CS = TracerAdvectCS()
    CS.diag = diag
    if "DT" not in param_file:   # fail_if_missing=.true. :1113
        raise MOMFatalError("MOM_tracer_advect: DT is required (fail_if_missing)")
    s_to_T = getattr(US, "s_to_T", 1.0) if US is not None else 1.0
    CS.dt = float(param_file["DT"]) * s_to_T
    CS.debug = bool(param_file.get("DEBUG", False))
    mesg = str(param_file.get("TRACER_ADVECTION_SCHEME", "PLM")).strip()
    if mesg == "PLM":            # :1122-1134
        CS.usePPM = False
    elif mesg == "PPM:H3":
        CS.usePPM = True
        CS.useHuynh = True
    elif mesg == "PPM":
        CS.usePPM = True
        CS.useHuynh = False
    else:
        raise MOMFatalError("MOM_tracer_advect, tracer_advect_init: Unknown TRACER_ADVECTION_SCHEME = " + mesg)
    if CS.useHuynh:              # :1136-1144
        CS.useHuynhStencilBug = bool(param_file.get("USE_HUYNH_STENCIL_BUG", False))
    return CS


def tracer_advect_end(CS: Optional[TracerAdvectCS]) -> None:
    """tracer_advect_end, MOM_tracer_advect.F90:1153-1158."""
    if CS is not None and CS._plan is not None:
        CS._plan.destroy()
        CS._plan = None


def max_iter_default(dt: float, CS: TracerAdvectCS) -> int:
    """:135  max_iter = 2*INT(CEILING(dt/CS%dt)) + 1"""
    return 2 * int(math.ceil(dt / CS.dt)) + 1


def _registry_signature(Reg):
    sig = [Reg.ntr]
    for tr in Reg.Tr[:Reg.ntr]:
        sig.append((tr.t.ctypes.data if isinstance(tr.t, np.ndarray) else tr.t.data_ptr(), tr.conc_underflow) +
                   tuple(None if getattr(tr, n) is None else id(getattr(tr, n))
                         for n in ("ad_x", "ad_y", "ad2d_x", "ad2d_y", "advection_xy")))
    return tuple(sig)


def _ensure_plan(G, GV, CS, Reg, device: int = -1):
    """Create the GPU plan on first use and (re)bind the registry when its
    pointers or ntr changed ('cheap hash check in the shim')."""
    from . import capi
    plan = CS._plan
    if plan is None:
        plan = capi.Plan(G, GV, CS, device=device)
        for k, v in CS.gpu_options.items():
            plan.set_option(k, int(v))
        if G.Domain.layout[0] * G.Domain.layout[1] > 1:     # LAYOUT > 1x1: exchange lists + NCCL communicator
            from . import multi
            multi.comm_init_plan(plan, G)
        plan._sig = None
        CS._plan = plan
    sig = _registry_signature(Reg)
    if plan._sig != sig:
        plan.set_ntr(Reg.ntr)
        for m, tr in enumerate(Reg.Tr[:Reg.ntr]):
            plan.bind_tracer(m, tr.t, tr.conc_underflow, tr.ad_x, tr.ad_y, tr.ad2d_x, tr.ad2d_y, tr.advection_xy)
        plan._sig = sig
    return plan


def advect_tracer(h_end, uhtr, vhtr, OBC, dt, G, GV, US, CS, Reg, x_first_in=None, vol_prev=None,
                  max_iter_in=None, update_vol_prev=None, uhr_out=None, vhr_out=None) -> int:
    """advect_tracer, MOM_tracer_advect.F90:52-324 -- same argument list.

    MIRROR mode of the C ABI: every array is a host (numpy) or device (torch)
    array in the reference's layout; Reg%Tr(m)%t is updated in place, associated
    diagnostics are filled, optional outputs are written.  Returns the number of
    iterations taken (the reference returns nothing)."""
    if CS is None:    # :120-121
        raise MOMFatalError("MOM_tracer_advect: tracer_advect_init must be called before advect_tracer.")
    if Reg is None:   # :122-123
        raise MOMFatalError("MOM_tracer_advect: register_tracer must be called before advect_tracer.")
    if Reg.ntr == 0:  # :124
        return 0
    from . import capi
    plan = _ensure_plan(G, GV, CS, Reg)
    try:
        if OBC is not None or getattr(plan, "_obc_set", False):    # reservoirs evolve: handed over on every call
            plan.set_obc(OBC)
            plan._obc_set = OBC is not None
        return plan.advect(h_end, uhtr, vhtr, dt,
                           x_first=-1 if x_first_in is None else int(bool(x_first_in)),
                           vol_prev=vol_prev,
                           max_iter_in=-1 if max_iter_in is None else int(max_iter_in),
                           update_vol_prev=-1 if update_vol_prev is None else int(bool(update_vol_prev)),
                           uhr_out=uhr_out, vhr_out=vhr_out)
    except capi.Mom6AdvError as e:
        raise MOMFatalError(str(e)) from e
