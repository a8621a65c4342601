"""Host-side mirror of the piece of offline_advection_ale (/root/reference/src/tracer/MOM_offline_main.F90:206-425)
that runs on the device: the Python twin of `offline_advection_iterate` in fortran/MOM_tracer_advect.F90
(INTEGRATION.md 4d).  h_pre, uhtr, vhtr go up once and come back once; the tracers stay in HBM."""
import numpy as np


def offline_advection_iterate(CS, G, GV, h_pre, uhtr, vhtr, dt_offline, num_off_iter, off_ale_mod, x_before_y,
                              min_residual, H_to_kg_m2=1.0, L_to_m=1.0, kg_m2_to_H=1.0, m_to_L=1.0):
    """Lines 298-415 of offline_advection_ale for runs that keep ALE out of the loop.  CS is the tracer_advect control
    structure that owns the GPU plan (made by an advect_tracer call or by the caller); fails loudly when there is none.
    h_pre, uhtr, vhtr (numpy, Fortran shapes [k, j, i]) are updated in place.  Returns (converged, trips, residuals)."""
    from . import capi
    plan = getattr(CS, "_plan", None)
    if plan is None:
        raise RuntimeError("offline_advection_iterate: the GPU plan does not exist yet (there is no CPU fallback)")
    plan.upload_field(capi.F_H_END, h_pre)
    plan.upload_field(capi.F_UHTR, uhtr)
    plan.upload_field(capi.F_VHTR, vhtr)
    out = plan.offline_advect(dt_offline, num_off_iter, x_before_y=x_before_y, off_ale_mod=off_ale_mod,
                              min_residual=min_residual, H_to_kg_m2=H_to_kg_m2, L_to_m=L_to_m, kg_m2_to_H=kg_m2_to_H,
                              m_to_L=m_to_L)
    plan.download_field(capi.F_H_END, h_pre)
    plan.download_field(capi.F_UHR, uhtr)
    plan.download_field(capi.F_VHR, vhtr)
    return out
