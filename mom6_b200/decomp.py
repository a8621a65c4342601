"""2-D block decomposition of the horizontal domain (MOM6 LAYOUT = pi,pj).

Mirrors what MOM_domains_init / create_MOM_domain set up
(src/framework/MOM_domains.F90:66, config_src/infra/FMS2/MOM_domain_infra.F90:1313):
each rank owns an (ni/pi) x (nj/pj) block with the same halo width; local
indexing restarts at isd = 1 (MOM_domain_infra.F90:1815-1821).
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np

from .types import MOMDomain, OceanGrid


def compute_extent(n: int, ndiv: int) -> List[Tuple[int, int]]:
    """Even split of 1..n into ndiv contiguous pieces (larger pieces in the
    middle is FMS's choice; any contiguous split is layout-reproducible, and we
    put the remainder on the first pieces)."""
    base, rem = divmod(n, ndiv)
    out, s = [], 1
    for p in range(ndiv):
        sz = base + (1 if p < rem else 0)
        out.append((s, s + sz - 1))
        s += sz
    return out


def rank_of(pi: int, pj: int, layout) -> int:
    return pi + layout[0] * pj


def local_grid(Gg: OceanGrid, layout, pe) -> OceanGrid:
    """Local OceanGrid of block ``pe`` = (pi, pj) cut from the global grid Gg."""
    npi, npj = layout
    Dg = Gg.Domain
    halo_i, halo_j = Gg.isc - Gg.isd, Gg.jsc - Gg.jsd
    (igs, ige) = compute_extent(Dg.nig, npi)[pe[0]]
    (jgs, jge) = compute_extent(Dg.njg, npj)[pe[1]]
    dom = MOMDomain(Dg.nig, Dg.njg, layout=tuple(layout), cyclic_x=Dg.cyclic_x, cyclic_y=Dg.cyclic_y,
                    tripolar_n=Dg.tripolar_n, symmetric=Dg.symmetric, pe=tuple(pe))
    G = OceanGrid(ige - igs + 1, jge - jgs + 1, halo=halo_i, jhalo=halo_j, symmetric=Gg.symmetric,
                  first_direction=Gg.first_direction, domain=dom)
    G.ioff = igs - G.isc     # global = local + off
    G.joff = jgs - G.jsc
    for name in ("areaT", "IareaT", "mask2dT"):
        setattr(G, name, cut_h(Gg, G, getattr(Gg, name)))
    G.mask2dCu = cut_u(Gg, G, Gg.mask2dCu)
    G.mask2dCv = cut_v(Gg, G, Gg.mask2dCv)
    return G


def _cols(Gg, G, lo_local, n, lo_global_array):
    # python column indices into the global array for local columns lo_local .. lo_local+n-1
    g = np.arange(n) + lo_local + G.ioff          # global (unwrapped) index of each local column
    gg = g - Gg.ioff                               # the global grid's own local index
    return gg - lo_global_array


def cut_h(Gg: OceanGrid, G: OceanGrid, a, index_only: bool = False):
    """Local data-domain copy of a global h-point array (halo included)."""
    ci = np.arange(G.nid) + G.isd + G.ioff - Gg.ioff - Gg.isd
    rj = np.arange(G.njd) + G.jsd + G.joff - Gg.joff - Gg.jsd
    if index_only:
        return rj, ci
    return np.ascontiguousarray(a[..., rj[:, None], ci[None, :]])


def cut_u(Gg, G, a, index_only: bool = False):
    ci = np.arange(G.nIB) + G.IsdB + G.ioff - Gg.ioff - Gg.IsdB
    rj = np.arange(G.njd) + G.jsd + G.joff - Gg.joff - Gg.jsd
    if index_only:
        return rj, ci
    return np.ascontiguousarray(a[..., rj[:, None], ci[None, :]])


def cut_v(Gg, G, a, index_only: bool = False):
    ci = np.arange(G.nid) + G.isd + G.ioff - Gg.ioff - Gg.isd
    rj = np.arange(G.nJB) + G.JsdB + G.joff - Gg.joff - Gg.JsdB
    if index_only:
        return rj, ci
    return np.ascontiguousarray(a[..., rj[:, None], ci[None, :]])


def paste_interior(Gg: OceanGrid, G: OceanGrid, a_global: np.ndarray, a_local: np.ndarray) -> None:
    """Write the computational domain of a local h-point array into the global one."""
    gi0 = G.isc + G.ioff - Gg.ioff - Gg.isd
    gj0 = G.jsc + G.joff - Gg.joff - Gg.jsd
    a_global[..., gj0:gj0 + G.nj, gi0:gi0 + G.ni] = G.interior(a_local)


def cut_obc(OBC, Gg: OceanGrid, G: OceanGrid):
    """The ocean_OBC_type a PE holds for the global open boundaries OBC: every segment's index ranges moved into local
    index space and clipped to the PE's data domain as setup_segment_indices does
    (/root/reference/src/core/MOM_open_boundary.F90:1296-1328), `on_pe` by the rule of setup_u_point_obc /
    setup_v_point_obc (:1443-1447: the face at least two points inside the data domain, the range overlapping it), the
    reservoirs cut to the local segment shape.  A segment that is not on the PE keeps its place in the list with
    is_E_or_W = is_N_or_S = False and no registry (:1446 returns before they are set), so segment numbers agree on all
    PEs.  OBC_pe is true where any segment is on the PE; the *_exist_globally flags are global."""
    from .types import OceanOBC, OBCSegment, OBCSegmentTracer, SegmentHI
    if OBC is None:
        return None
    di, dj = G.ioff - Gg.ioff, G.joff - Gg.joff        # local = global-grid index - d
    clip = lambda v, lo, hi: min(max(v, lo), hi)
    segs, any_on = [], False
    for sg in OBC.segment:
        H = sg.HI
        hi = SegmentHI(isd=clip(H.isd - di, G.isd, G.ied), ied=clip(H.ied - di, G.isd, G.ied),
                       jsd=clip(H.jsd - dj, G.jsd, G.jed), jed=clip(H.jed - dj, G.jsd, G.jed),
                       IsdB=clip(H.IsdB - di, G.IsdB, G.IedB), IedB=clip(H.IedB - di, G.IsdB, G.IedB),
                       JsdB=clip(H.JsdB - dj, G.JsdB, G.JedB), JedB=clip(H.JedB - dj, G.JsdB, G.JedB))
        if sg.is_E_or_W:
            I_obc, Js_obc, Je_obc = H.IsdB - di, H.JsdB - dj, H.JedB - dj
            on = not (I_obc <= G.IsdB + 1 or I_obc >= G.IedB - 1) and not (Je_obc <= G.JsdB or Js_obc >= G.JedB)
        else:
            J_obc, Is_obc, Ie_obc = H.JsdB - dj, H.IsdB - di, H.IedB - di
            on = not (J_obc <= G.JsdB + 1 or J_obc >= G.JedB - 1) and not (Ie_obc <= G.IsdB or Is_obc >= G.IedB)
        if not on:
            segs.append(OBCSegment(direction=sg.direction, specified=sg.specified, HI=hi, tr_Reg=None))
            continue
        any_on = True
        reg = None
        if sg.tr_Reg is not None:
            reg = []
            for t in sg.tr_Reg:
                tres = None
                if t.tres is not None:
                    if sg.is_E_or_W:      # [k, j, I] over (IsdB:IedB, jsd:jed)
                        tres = np.ascontiguousarray(t.tres[:, hi.jsd + dj - H.jsd:hi.jed + dj - H.jsd + 1, hi.IsdB + di - H.IsdB:hi.IedB + di - H.IsdB + 1])
                    else:                 # [k, J, i] over (isd:ied, JsdB:JedB)
                        tres = np.ascontiguousarray(t.tres[:, hi.JsdB + dj - H.JsdB:hi.JedB + dj - H.JsdB + 1, hi.isd + di - H.isd:hi.ied + di - H.isd + 1])
                reg.append(OBCSegmentTracer(ntr_index=t.ntr_index, tres=tres, OBC_inflow_conc=t.OBC_inflow_conc))
        segs.append(OBCSegment(direction=sg.direction, is_E_or_W=sg.is_E_or_W, is_N_or_S=sg.is_N_or_S, specified=sg.specified,
                               HI=hi, tr_Reg=reg))
    return OceanOBC(OBC_pe=any_on, segment=segs, open_u_BCs_exist_globally=OBC.open_u_BCs_exist_globally,
                    open_v_BCs_exist_globally=OBC.open_v_BCs_exist_globally,
                    specified_u_BCs_exist_globally=OBC.specified_u_BCs_exist_globally,
                    specified_v_BCs_exist_globally=OBC.specified_v_BCs_exist_globally)
