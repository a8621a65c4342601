"""Host-side mirror of the pieces of tracer_hordiff (/root/reference/src/tracer/MOM_tracer_hor_diff.F90) that sit
around the device call: the iteration count from the diffusive CFL (:368-387) and the replacement of the iteration
loop (:539-607) -- the Python twin of `tracer_hordiff_iterate` in fortran/MOM_tracer_advect.F90.
The 2-D diffusivity set-up (:186-366) stays with the caller, as in the Fortran integration."""
import math

import numpy as np


def _max_across_pes(x: float) -> float:
    """max_across_PEs (:377): torch.distributed MAX over the ranks when a process group with more than one rank exists."""
    try:
        import torch
        import torch.distributed as dist
    except Exception:
        return x
    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return x
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def num_itts_from_cfl(khdt_x, khdt_y, G, check_diffusive_CFL=True, max_diff_CFL=-1.0) -> int:
    """num_itts as :368-387 computes it from this rank's block, max_across_PEs (:377) included: one rank per GPU with
    torch.distributed initialised gives every rank the global value; a single process gets its own."""
    if check_diffusive_CFL:
        kx, ky, ia = khdt_x, khdt_y, G.IareaT
        j0, j1, i0, i1 = G.pj(G.jsc), G.pj(G.jec), G.pi(G.isc), G.pi(G.iec)
        J0, I0 = G.pJ(G.jsc), G.pI(G.isc)
        nj, ni = j1 - j0 + 1, i1 - i0 + 1
        cfl = 2.0 * ((kx[j0:j1 + 1, I0 - 1:I0 - 1 + ni] + kx[j0:j1 + 1, I0:I0 + ni]) +
                     (ky[J0 - 1:J0 - 1 + nj, i0:i1 + 1] + ky[J0:J0 + nj, i0:i1 + 1])) * ia[j0:j1 + 1, i0:i1 + 1]
        mx = _max_across_pes(max(0.0, float(np.asarray(cfl).max())))
        return max(1, math.ceil(mx - 4.0 * np.finfo(np.float64).eps))
    if max_diff_CFL > 0.0:
        return max(1, math.ceil(max_diff_CFL - 4.0 * np.finfo(np.float64).eps))
    return 1


def tracer_hordiff_iterate(CS, G, GV, khdt_x, khdt_y, num_itts, Diffuse_ML_interior=False, ML_KhTr_scale=0.0,
                           any_df=False, nkml=None, nk_rho_varies=None) -> None:
    """The loop `do itt=1,num_itts ... enddo` (:539-607) on the tracers the last advect_tracer call left in HBM.
    CS is the tracer_advect control structure that owns the GPU plan; fails loudly when there is none."""
    plan = getattr(CS, "_plan", None)
    if plan is None:
        raise RuntimeError("tracer_hordiff_iterate: no advect_tracer call has made the GPU plan yet (there is no CPU fallback)")
    plan.hordiff(khdt_x, khdt_y, int(num_itts), GV.H_subroundoff, h=None, Diffuse_ML_interior=Diffuse_ML_interior,
                 nkml=int(getattr(GV, "nkml", 0) if nkml is None else nkml),
                 nk_rho_varies=int(getattr(GV, "nk_rho_varies", 0) if nk_rho_varies is None else nk_rho_varies),
                 ML_KhTr_scale=ML_KhTr_scale, want_df=any_df)
