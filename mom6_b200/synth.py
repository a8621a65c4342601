"""Seeded synthetic inputs of the BASELINE.json shapes.

Everything is generated with torch float64 on ``device`` (cpu for parity tests,
where identical bytes go to the oracle and to the GPU; cuda for full-size
bench runs, where no host copy of 48 GB of tracers ever exists).

Grid: Mercator-like ``areaT = A0*cos(lat_j)``, lat in [-78, 65]; ~30 % land from a
thresholded smooth field; cyclic in x (optional), closed south and north.
Thickness: nk-level z*-like profile with partial/vanished layers at the
synthetic bathymetry (Angstrom_H below it and on land).  Transports: per-layer
corner streamfunction (gyres + mesoscale modes) plus a 10 % divergent part,
masked by mask2dCu/Cv, scaled to a target upwind-CFL.  h_end is made consistent
with the transports so that the reconstructed hprev is ~ areaT*h_start > 0.
Tracers: T (smooth + sharp fronts), S, then passive tracers cycling through
the advection_test_tracer shapes (src/tracer/advection_test_tracer.F90:206-238:
square, triangle, cosine bell, cylinder, cut cylinder) and band-limited noise.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from .types import MOMDomain, OceanGrid, VerticalGrid

SEED0 = 0x4D4F4D36

# BASELINE.json configs: name -> (ni, nj, nk, ntr, scheme)
CONFIGS = {
    "tc1": (10, 8, 8, 3, "PLM"),
    "double_gyre": (44, 40, 2, 10, "PPM"),
    "OM4_1": (360, 210, 75, 22, "PPM"),
    "OM4_025": (1440, 1080, 75, 52, "PPM"),
    "eighth": (2880, 2160, 75, 22, "PPM"),
}


def _rng(seed: int) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64(seed))


@dataclass
class SynthCase:
    G: OceanGrid
    GV: VerticalGrid
    h_end: torch.Tensor      # [nk, njd, nid]
    uhtr: torch.Tensor       # [nk, njd, nIB]
    vhtr: torch.Tensor       # [nk, nJB, nid]
    ntr: int
    seed: int
    device: str
    dt: float = 3600.0
    dt_baroclinic: float = 900.0
    # cached coordinate helpers for the tracer generator
    _x: Optional[torch.Tensor] = None
    _y: Optional[torch.Tensor] = None
    _kk: Optional[torch.Tensor] = None
    _maskT: Optional[torch.Tensor] = None

    def conc_underflow(self, m: int) -> float:
        return 1.0e-35 if m < 2 else 0.0

    def tracer(self, m: int) -> torch.Tensor:
        return _gen_tracer(self, m)


def _smooth(x, y, rng, nmodes, kmax, periodic_x=True):
    """Band-limited field on the (x in [0,1), y in [0,1]) grids; exactly
    periodic in x only up to rounding, so callers copy wrap columns."""
    f = torch.zeros(torch.broadcast_shapes(x.shape, y.shape), dtype=torch.float64, device=x.device)
    for _ in range(nmodes):
        kx = int(rng.integers(1, kmax + 1)) if periodic_x else float(rng.uniform(0.5, kmax))
        ky = float(rng.uniform(0.5, kmax))
        ph1, ph2 = float(rng.uniform(0, 2 * math.pi)), float(rng.uniform(0, 2 * math.pi))
        amp = float(rng.uniform(0.5, 1.0)) / math.sqrt(kx * kx + ky * ky)
        f = f + amp * torch.sin(2 * math.pi * kx * x + ph1) * torch.sin(math.pi * ky * y + ph2)
    return f / (f.abs().max() + 1e-300)


def make_case(ni: int, nj: int, nk: int, ntr: int, halo: int = 4, seed: int = SEED0,
              device: str = "cpu", cyclic_x: bool = True, cyclic_y: bool = False,
              symmetric: bool = True, land: bool = True, cfl: float = 0.3,
              first_direction: int = 0, dt: float = 3600.0, dt_baroclinic: float = 900.0,
              tripolar_n: bool = False) -> SynthCase:
    """Build grid + h_end/uhtr/vhtr.  ``cfl`` is the approximate 99th-percentile
    upwind CFL of the transports (values near/above 0.5 engage the flux
    limiter and force residual iterations)."""
    dom = MOMDomain(ni, nj, cyclic_x=cyclic_x, cyclic_y=cyclic_y, symmetric=symmetric, tripolar_n=tripolar_n)
    G = OceanGrid(ni, nj, halo=halo, symmetric=symmetric, first_direction=first_direction, domain=dom)
    GV = VerticalGrid.default(nk)
    rng = _rng(seed)
    dev = torch.device(device)
    f64 = dict(dtype=torch.float64, device=dev)

    # global cell indices of every data-domain point (may lie outside 1..ni)
    ig = torch.arange(G.isd, G.ied + 1, **f64) + G.ioff            # [nid]
    jg = torch.arange(G.jsd, G.jed + 1, **f64) + G.joff            # [njd]
    xc = ((ig - 0.5) / ni).view(1, -1)
    yc = ((jg - 0.5) / nj).view(-1, 1)

    # ---- grid metrics ------------------------------------------------------
    lat = (-78.0 + 143.0 * yc).clamp(-89.0, 89.0) * (math.pi / 180.0)
    A0 = (4.0e7 / ni) * (4.0e7 * 143.0 / 360.0 / nj)
    areaT = (A0 * torch.cos(lat)).expand(G.njd, G.nid).contiguous()
    IareaT = 1.0 / areaT

    # ---- land mask ----------------------------------------------------------
    in_x = (ig >= 1) & (ig <= ni)
    in_y = (jg >= 1) & (jg <= nj)
    if land:
        lf = _smooth(xc, yc, rng, 6, 4)
        thr = torch.quantile(lf.flatten()[:: max(1, lf.numel() // 200000)], 0.3)
        maskT = (lf > thr).to(torch.float64)
    else:
        maskT = torch.ones(G.njd, G.nid, **f64)
    maskT = _wrap_h(G, dom, maskT, in_x, in_y, fill=0.0)
    mCu = torch.zeros(G.njd, G.nIB, **f64)
    # u face I (python col I-IsdB) sits between cells I and I+1
    c0 = G.pi(max(G.IsdB, G.isd)); I0 = G.pI(max(G.IsdB, G.isd))
    n = G.nid - 1 - c0
    mCu[:, I0:I0 + n] = maskT[:, c0:c0 + n] * maskT[:, c0 + 1:c0 + 1 + n]
    mCv = torch.zeros(G.nJB, G.nid, **f64)
    r0 = G.pj(max(G.JsdB, G.jsd)); J0 = G.pJ(max(G.JsdB, G.jsd))
    n = G.njd - 1 - r0
    mCv[J0:J0 + n, :] = maskT[r0:r0 + n, :] * maskT[r0 + 1:r0 + 1 + n, :]

    G.areaT, G.IareaT, G.mask2dT = _np(areaT), _np(IareaT), _np(maskT)
    G.mask2dCu, G.mask2dCv = _np(mCu), _np(mCv)

    # ---- thickness ----------------------------------------------------------
    dz = torch.tensor(_dz_profile(nk), **f64)                       # [nk]
    ztop = torch.cumsum(dz, 0) - dz
    depth_tot = float(dz.sum())
    bathy = depth_tot * (0.25 + 0.75 * (0.5 + 0.5 * _smooth(xc, yc, rng, 5, 3)))
    bathy = _wrap_h(G, dom, bathy * maskT, in_x, in_y, fill=0.0)
    noise = _wrap_h(G, dom, _smooth(xc, yc, rng, 8, 8), in_x, in_y, fill=0.0)
    ang = GV.Angstrom_H
    h_start = torch.empty(nk, G.njd, G.nid, **f64)
    for k in range(nk):
        hk = (bathy - ztop[k]).clamp(min=0.0, max=float(dz[k])) * (1.0 + 0.1 * noise * math.cos(0.37 * k))
        h_start[k] = torch.where(hk > ang, hk, torch.full_like(hk, ang))

    # ---- transports ---------------------------------------------------------
    # corner streamfunction at (I,J): x = Ig/ni, y = Jg/nj, python arrays [nJB', nIB'] over
    # Ig in isd-1..ied, Jg in jsd-1..jed (always including the 0 index)
    Ig = torch.arange(G.isd - 1, G.ied + 1, **f64) + G.ioff
    Jg = torch.arange(G.jsd - 1, G.jed + 1, **f64) + G.joff
    xq, yq = (Ig / ni).view(1, -1), (Jg / nj).view(-1, 1)
    gy = torch.sin(math.pi * yq.clamp(0, 1)) ** 2
    psi_g = gy * _smooth(xq, yq, rng, 3, 2)                         # gyres
    psi_m = gy * _smooth(xq, yq, rng, 12, max(4, min(ni, nj) // 6))  # mesoscale
    phi = _smooth(xc, yc, rng, 4, 3)                                # divergent potential (cells)
    uh = torch.zeros(nk, G.njd, G.nIB, **f64)
    vh = torch.zeros(nk, G.nJB, G.nid, **f64)
    vol = areaT.unsqueeze(0) * h_start                              # [nk, njd, nid]
    oI = 0 if symmetric else 1   # column offset of psi (starts at isd-1) relative to u arrays
    oJ = 0 if symmetric else 1
    gu = torch.zeros(G.njd, G.nid + 1, **f64); gu[:, 1:-1] = phi[:, 1:] - phi[:, :-1]
    gv = torch.zeros(G.njd + 1, G.nid, **f64); gv[1:-1, :] = phi[1:, :] - phi[:-1, :]

    def nondim(k):
        a = math.cos(0.21 * k)
        psi = (0.6 + 0.4 * a) * psi_g + 0.5 * math.sin(0.5 + 0.33 * k) * psi_m
        # uhtr(I,j) = psi(I,J) - psi(I,J-1): cell row j has corners J=j (north) and J-1 (south)
        du = (psi[1:, :] - psi[:-1, :]) + 0.1 * gu                  # [njd, nid+1] faces isd-1..ied
        dv = -(psi[:, 1:] - psi[:, :-1]) + 0.1 * gv                 # [njd+1, nid] faces jsd-1..jed
        return du, dv

    du0, dv0 = nondim(0)
    q = max(_q99(du0[:, oI:][mCu > 0]), _q99(dv0[oJ:, :][mCv > 0]), 1e-300)
    scale = cfl / q
    for k in range(nk):
        du, dv = nondim(k)
        # scale by the smaller neighbouring cell volume: |du|, |dv| bound the upwind CFL
        vu = torch.zeros_like(du); vu[:, 1:-1] = torch.minimum(vol[k][:, :-1], vol[k][:, 1:])
        vv = torch.zeros_like(dv); vv[1:-1, :] = torch.minimum(vol[k][:-1, :], vol[k][1:, :])
        uh[k] = (scale * du * vu)[:, oI:] * mCu
        vh[k] = (scale * dv * vv)[oJ:, :] * mCv
    # make the transports exactly consistent across the cyclic seam / zero outside the domain
    uh = _finish_u(G, dom, uh)
    vh = _finish_v(G, dom, vh)

    # ---- h_end consistent with the transports -------------------------------
    div = torch.zeros(nk, G.njd, G.nid, **f64)
    cI = G.pI(G.isd)      # python column in uh of face I == isd  (east face of cell isd)
    cJ = G.pJ(G.jsd)
    # cells isd+1..ied and jsd+1..jed have both faces inside the arrays
    div[:, 1:, 1:] = (uh[:, 1:, cI + 1:cI + G.nid] - uh[:, 1:, cI:cI + G.nid - 1]) + \
                     (vh[:, cJ + 1:cJ + G.njd, 1:] - vh[:, cJ:cJ + G.njd - 1, 1:])
    h_end = torch.maximum(h_start - div * IareaT.unsqueeze(0), torch.full_like(h_start, ang))

    case = SynthCase(G=G, GV=GV, h_end=h_end.contiguous(), uhtr=uh.contiguous(), vhtr=vh.contiguous(),
                     ntr=ntr, seed=seed, device=device, dt=dt, dt_baroclinic=dt_baroclinic)
    case._x, case._y, case._maskT = xc, yc, maskT
    return case


# ---------------------------------------------------------------------------
# helpers
# ---------------------------------------------------------------------------
def _np(t: torch.Tensor) -> np.ndarray:
    return np.ascontiguousarray(t.detach().cpu().numpy())


def _q99(x: torch.Tensor) -> float:
    x = x.abs().flatten()
    if x.numel() == 0:
        return 0.0
    x = x[:: max(1, x.numel() // 500000)]
    return float(torch.quantile(x, 0.99))


def _dz_profile(nk: int):
    """2 m at the surface growing to 250 m at depth (OM4-like 75 level z*)."""
    if nk == 1:
        return [100.0]
    return [2.0 * (125.0 ** (k / (nk - 1))) for k in range(nk)]


def _wrap_cols(G, dom, a, lo_index, n_global, fill):
    """Make the last axis consistent with the x topology.  ``lo_index`` is the
    global index of python column 0 (cells: isd+ioff)."""
    n = a.shape[-1]
    gidx = torch.arange(n, device=a.device) + lo_index
    outside = (gidx < 1) | (gidx > n_global)
    if not bool(outside.any()):
        return a
    if dom.cyclic_x:
        src = ((gidx - 1) % n_global) + 1 - lo_index
        return a.index_select(-1, src)
    a = a.clone()
    a[..., outside] = fill
    return a


def _wrap_rows(G, dom, a, lo_index, n_global, fill, fold=False):
    n = a.shape[-2]
    gidx = torch.arange(n, device=a.device) + lo_index
    outside = (gidx < 1) | (gidx > n_global)
    if not bool(outside.any()):
        return a
    if dom.cyclic_y:
        src = ((gidx - 1) % n_global) + 1 - lo_index
        return a.index_select(-2, src)
    a = a.clone()
    a[..., outside, :] = fill
    if fold and dom.tripolar_n:
        # tripolar north fold of an h-point field (src/framework/MOM_domains.F90:184-191; FMS FOLD_NORTH_EDGE): halo row
        # njg + d holds row njg + 1 - d with the columns reversed, ig -> nig + 1 - ig (columns are already x-wrapped)
        ncol = a.shape[-1]
        ig = torch.arange(ncol, device=a.device) + (G.isd + G.ioff)
        csrc = (dom.nig + 1 - ig) - (G.isd + G.ioff)
        ok = (csrc >= 0) & (csrc < ncol)
        for r in torch.nonzero(gidx > n_global).flatten().tolist():
            rs = (2 * n_global + 1 - int(gidx[r])) - lo_index
            if 0 <= rs < n:
                row = a[..., rs, :].index_select(-1, csrc.clamp(0, ncol - 1))
                a[..., r, :] = torch.where(ok, row, torch.full_like(row, fill))
    return a


def _wrap_h(G, dom, a, in_x=None, in_y=None, fill=0.0):
    a = _wrap_cols(G, dom, a, G.isd + G.ioff, dom.nig, fill)
    return _wrap_rows(G, dom, a, G.jsd + G.joff, dom.njg, fill, fold=True)


def _finish_u(G, dom, uh):
    # u faces: global index Ig = I + ioff in 0..nig; canonical range 1..nig (0 == nig when cyclic)
    n = uh.shape[-1]
    gidx = torch.arange(n, device=uh.device) + (G.IsdB + G.ioff)
    if dom.cyclic_x:
        src = ((gidx - 1) % dom.nig) + 1 - (G.IsdB + G.ioff)
        uh = uh.index_select(-1, src)
    else:
        uh = uh.clone()
        uh[..., (gidx < 0) | (gidx > dom.nig)] = 0.0
    return _wrap_rows(G, dom, uh, G.jsd + G.joff, dom.njg, 0.0)


def _finish_v(G, dom, vh):
    n = vh.shape[-2]
    gidx = torch.arange(n, device=vh.device) + (G.JsdB + G.joff)
    if dom.cyclic_y:
        src = ((gidx - 1) % dom.njg) + 1 - (G.JsdB + G.joff)
        vh = vh.index_select(-2, src)
    else:
        vh = vh.clone()
        vh[..., (gidx < 0) | (gidx > dom.njg), :] = 0.0
    return _wrap_cols(G, dom, vh, G.isd + G.ioff, dom.nig, 0.0)


# ---------------------------------------------------------------------------
# tracers
# ---------------------------------------------------------------------------
def _gen_tracer(case: SynthCase, m: int) -> torch.Tensor:
    """Tracer m as a [nk, njd, nid] float64 tensor on case.device.  Land cells
    and closed-boundary halos hold finite junk (they must never matter)."""
    G, nk = case.G, case.GV.ke
    x, y, maskT = case._x, case._y, case._maskT
    dev = x.device
    rng = _rng(case.seed * 1000003 + 7919 * (m + 1))
    kk = torch.arange(nk, dtype=torch.float64, device=dev).view(-1, 1, 1) / max(nk - 1, 1)
    x3, y3 = x.unsqueeze(0), y.unsqueeze(0)
    if m == 0:      # temperature: smooth + two sharp fronts, about -2..30
        f = 14.0 + 12.0 * torch.cos(math.pi * (y3 - 0.45)) * (1.0 - 0.7 * kk)
        f = f + 3.0 * torch.tanh((y3 - 0.3 - 0.05 * torch.sin(2 * math.pi * 3 * x3)) * 80.0)
        f = f - 4.0 * torch.tanh((y3 - 0.7 - 0.04 * torch.cos(2 * math.pi * 5 * x3)) * 120.0)
        f = f + 0.5 * _smooth(x, y, rng, 6, 6).unsqueeze(0)
    elif m == 1:    # salinity 33..37
        f = 35.0 + 1.5 * _smooth(x, y, rng, 6, 5).unsqueeze(0) * (1.0 - 0.5 * kk) + 0.3 * kk
    else:
        shape = (m - 2) % 6
        x0, y0 = float(rng.uniform(0.1, 0.9)), float(rng.uniform(0.2, 0.8))
        L = float(rng.uniform(0.08, 0.2))
        dx = torch.remainder(x3 - x0 + 0.5, 1.0) - 0.5
        dy = y3 - y0 - 0.1 * (kk - 0.5)
        r = torch.sqrt(dx * dx + dy * dy) / L
        if shape == 0:    # square
            f = ((dx.abs() < L) & (dy.abs() < L)).to(torch.float64)
        elif shape == 1:  # triangle
            f = (1.0 - dx.abs() / L).clamp(min=0.0) * (dy.abs() < L).to(torch.float64)
        elif shape == 2:  # cosine bell
            f = torch.where(r < 1.0, 0.5 * (1.0 + torch.cos(math.pi * r)), torch.zeros_like(r))
        elif shape == 3:  # cylinder
            f = (r < 1.0).to(torch.float64)
        elif shape == 4:  # cut cylinder
            f = ((r < 1.0) & ~((dx.abs() < 0.25 * L) & (dy > -0.5 * L))).to(torch.float64)
        else:             # band-limited noise in [0,1]
            f = 0.5 + 0.5 * _smooth(x, y, rng, 10, 10).unsqueeze(0) * torch.cos(2.0 * kk + m)
        f = f * (1.0 + 0.01 * m) + 0.0 * kk
    f = f.expand(nk, G.njd, G.nid)
    junk = 1234.5 + 0.25 * m
    out = torch.where(maskT.unsqueeze(0) > 0.5, f, torch.full((1, 1, 1), junk, dtype=torch.float64, device=dev))
    return out.contiguous()


def bind_registry(case: SynthCase, diags: bool = False):
    """Host (numpy) tracer registry for a CPU-generated case."""
    from .types import TracerRegistry, register_tracer
    Reg = TracerRegistry()
    G, nk = case.G, case.GV.ke
    for m in range(case.ntr):
        t = _np(case.tracer(m))
        tr = register_tracer(t, Reg, name="tr%02d" % m, underflow_conc=case.conc_underflow(m))
        if diags:
            tr.ad_x = np.full(G.shape_u(nk), np.nan)
            tr.ad_y = np.full(G.shape_v(nk), np.nan)
            tr.ad2d_x = np.full(G.shape_u(), np.nan)
            tr.ad2d_y = np.full(G.shape_v(), np.nan)
            tr.advection_xy = np.full(G.shape_h(nk), np.nan)
    return Reg


def make_obc_case(ni: int = 40, nj: int = 34, nk: int = 3, ntr: int = 5, cfl: float = 0.6, seed: int = SEED0 + 77,
                  halo: int = 4, symmetric: bool = True):
    """A closed rectangular basin whose outer ring of cells is masked, with one open-boundary segment on each
    side (regional-model geometry, cf. the DOME tracers of the reference's tc3):

      W  specified, tracer reservoirs tres for tracers 1..3          face I = isc,   rows jsc+2 .. jec-2
      E  open (radiation-like), OBC_inflow_conc only, tracers 2,4    face I = iec-1, rows jsc+1 .. jec-3
      S  specified, tres for tracer 1, inflow_conc for tracer 5      face J = jsc,   columns isc+2 .. iec-2
      N  open, tres for tracers 1..ntr                               face J = jec-1, columns isc+3 .. iec-1
      +  one E/W segment whose tr_Reg is unassociated (must be skipped)

    Transports through the segment faces have both signs, so inflow (prescribed flux, uhh := uhr) and outflow
    (interior scheme with the overridden exterior value) both occur.  Returns (case, OBC)."""
    from .types import (OceanOBC, OBCSegment, OBCSegmentTracer, SegmentHI, OBC_DIRECTION_N, OBC_DIRECTION_S,
                        OBC_DIRECTION_E, OBC_DIRECTION_W)
    c = make_case(ni, nj, nk, ntr, halo=halo, seed=seed, cyclic_x=False, land=False, cfl=cfl, symmetric=symmetric)
    G = c.G
    rng = _rng(seed + 1)
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    # masked outer ring
    mT = G.mask2dT
    mT[...] = 0.0
    mT[G.pj(G.jsc + 1):G.pj(G.jec), G.pi(G.isc + 1):G.pi(G.iec)] = 1.0
    G.mask2dCu[...] = 0.0
    G.mask2dCv[...] = 0.0
    for j in range(G.jsd, G.jed + 1):
        for I in range(max(G.IsdB, G.isd), G.ied):
            G.mask2dCu[G.pj(j), G.pI(I)] = mT[G.pj(j), G.pi(I)] * mT[G.pj(j), G.pi(I + 1)]
    for J in range(max(G.JsdB, G.jsd), G.jed):
        for i in range(G.isd, G.ied + 1):
            G.mask2dCv[G.pJ(J), G.pi(i)] = mT[G.pj(J), G.pi(i)] * mT[G.pj(J + 1), G.pi(i)]
    # exterior cells of the western side are thin (Angstrom), as masked columns are in the model: an inflow then
    # finds an empty upwind cell, the limiter returns uhh = 0 and the OBC branch restores uhh = uhr (:592-600)
    h[:, :, G.pi(G.isc)] = c.GV.Angstrom_H
    u *= G.mask2dCu[None]
    v *= G.mask2dCv[None]
    amp_u = float(np.abs(u).max()) or 1.0
    amp_v = float(np.abs(v).max()) or 1.0

    def seg_ew(I, j0, j1, direction, specified, trs):
        hi = SegmentHI(isd=I, ied=I + 1, jsd=j0, jed=j1, IsdB=I, IedB=I, JsdB=j0 - 1, JedB=j1)
        G.mask2dCu[G.pj(j0):G.pj(j1) + 1, G.pI(I)] = 1.0            # the open face
        u[:, G.pj(j0):G.pj(j1) + 1, G.pI(I)] = amp_u * rng.uniform(-1.0, 1.0, (nk, j1 - j0 + 1))
        reg = None
        if trs is not None:
            reg = []
            for (m1, use_tres, conc) in trs:
                tres = np.ascontiguousarray(rng.uniform(0.0, 30.0, (nk, j1 - j0 + 1, 1))) if use_tres else None
                reg.append(OBCSegmentTracer(ntr_index=m1, tres=tres, OBC_inflow_conc=conc))
        return OBCSegment(direction=direction, is_E_or_W=True, specified=specified, HI=hi, tr_Reg=reg)

    def seg_ns(J, i0, i1, direction, specified, trs):
        hi = SegmentHI(isd=i0, ied=i1, jsd=J, jed=J + 1, IsdB=i0 - 1, IedB=i1, JsdB=J, JedB=J)
        G.mask2dCv[G.pJ(J), G.pi(i0):G.pi(i1) + 1] = 1.0
        v[:, G.pJ(J), G.pi(i0):G.pi(i1) + 1] = amp_v * rng.uniform(-1.0, 1.0, (nk, i1 - i0 + 1))
        reg = []
        for (m1, use_tres, conc) in trs:
            tres = np.ascontiguousarray(rng.uniform(0.0, 30.0, (nk, 1, i1 - i0 + 1))) if use_tres else None
            reg.append(OBCSegmentTracer(ntr_index=m1, tres=tres, OBC_inflow_conc=conc))
        return OBCSegment(direction=direction, is_N_or_S=True, specified=specified, HI=hi, tr_Reg=reg)

    nt = min(ntr, 3)
    segs = [
        seg_ew(G.isc, G.jsc + 2, G.jec - 2, OBC_DIRECTION_W, True, [(m + 1, True, 0.0) for m in range(nt)]),
        seg_ew(G.iec - 1, G.jsc + 1, G.jec - 3, OBC_DIRECTION_E, False, [(min(2, ntr), False, 7.5), (min(4, ntr), False, -1.25)]),
        seg_ns(G.jsc, G.isc + 2, G.iec - 2, OBC_DIRECTION_S, True, [(1, True, 0.0), (ntr, False, 3.0)]),
        seg_ns(G.jec - 1, G.isc + 3, G.iec - 1, OBC_DIRECTION_N, False, [(m + 1, True, 0.0) for m in range(ntr)]),
        seg_ew(G.isc + 6, G.jsc + 4, G.jsc + 8, OBC_DIRECTION_W, False, None),
    ]
    # the last segment is an interior dummy: undo its face edits (its tr_Reg is unassociated, it must be ignored)
    I = G.isc + 6
    for j in range(G.jsc + 4, G.jsc + 9):
        G.mask2dCu[G.pj(j), G.pI(I)] = mT[G.pj(j), G.pi(I)] * mT[G.pj(j), G.pi(I + 1)]
    OBC = OceanOBC(OBC_pe=True, segment=segs, open_u_BCs_exist_globally=True, open_v_BCs_exist_globally=True,
                   specified_u_BCs_exist_globally=True, specified_v_BCs_exist_globally=True)
    return c, OBC


def transposed_case(c: "SynthCase", Reg):
    """The same problem with i and j exchanged (the reference's index-rotation standard, .testing/Makefile:384,455,596, in
    its simplest form): grid metrics and masks transposed, uhtr <-> vhtr, REENTRANT_X <-> REENTRANT_Y, every tracer
    transposed.  Advecting it with the sweep order flipped must give the transposed answer.  Symmetric memory only.
    Returns (G2, h2, u2, v2, Reg2) with numpy arrays."""
    from .types import TracerRegistry, register_tracer
    G, D = c.G, c.G.Domain
    assert G.symmetric, "transposition needs symmetric memory (u and v arrays of matching shapes)"
    D2 = MOMDomain(D.njg, D.nig, cyclic_x=D.cyclic_y, cyclic_y=D.cyclic_x, symmetric=True)
    G2 = OceanGrid(G.nj, G.ni, halo=G.jsc - G.jsd, jhalo=G.isc - G.isd, symmetric=True,
                   first_direction=G.first_direction + 1, domain=D2)
    tr2 = lambda a: np.ascontiguousarray(np.swapaxes(np.asarray(a), -1, -2))
    G2.areaT, G2.IareaT, G2.mask2dT = tr2(G.areaT), tr2(G.IareaT), tr2(G.mask2dT)
    G2.mask2dCu, G2.mask2dCv = tr2(G.mask2dCv), tr2(G.mask2dCu)
    h2, u2, v2 = tr2(c.h_end.cpu().numpy()), tr2(c.vhtr.cpu().numpy()), tr2(c.uhtr.cpu().numpy())
    Reg2 = TracerRegistry()
    for tr in Reg.Tr[:Reg.ntr]:
        register_tracer(tr2(tr.t), Reg2, name=getattr(tr, "name", "tr"), underflow_conc=tr.conc_underflow)
    return G2, h2, u2, v2, Reg2


def make_khdt(c: "SynthCase", strength: float = 0.3, seed: int = 11):
    """khdt_x (jsd:jed, IsdB:IedB) and khdt_y (JsdB:JedB, isd:ied) as tracer_hordiff's 2-D set-up would leave them
    (MOM_tracer_hor_diff.F90:273-366): Khtr*dt*(open face width / tracer-point distance), zero at closed faces, with
    valid halos.  `strength` ~ khdt / areaT: the diffusive CFL 2*(sum of the four khdt)/areaT is about 8*strength, so
    0.3 asks for three iterations (:368-387)."""
    G = c.G
    rng = np.random.default_rng(seed)
    a = np.asarray(G.areaT)
    kx = np.zeros(G.shape_u()); ky = np.zeros(G.shape_v())
    oI = G.pI(G.isd)      # u column of face I = isd (between cells isd and isd+1)
    n = G.nid - 1
    kx[:, oI:oI + n] = strength * np.minimum(a[:, :-1], a[:, 1:]) * (0.6 + 0.4 * rng.random((G.njd, n)))
    oJ = G.pJ(G.jsd)
    n = G.njd - 1
    ky[oJ:oJ + n, :] = strength * np.minimum(a[:-1, :], a[1:, :]) * (0.6 + 0.4 * rng.random((n, G.nid)))
    return np.ascontiguousarray(kx * np.asarray(G.mask2dCu)), np.ascontiguousarray(ky * np.asarray(G.mask2dCv))
