// x sweep of the streaming path (see fast_common.cuh for the design notes).
#include "fast_common.cuh"
#include <algorithm>

// ---------------------------------------------------------------------------
// x sweep
// ---------------------------------------------------------------------------
#define XT_USE 120     // useful cells per 128-cell warp tile (lanes 1..30)

// Rows or warp tiles in which no cell is updated (:636-649 do_i false everywhere: land, vanished transports):
// copy through, with the underflow test on the sweep's own range (:683-687).
__device__ __forceinline__ void x_copy_through(const TrTable &tt, long long q0, bool st, const bool (&inr)[4]) {
  if (!st) return;
  // four tracers per round so that eight 16-byte loads are in flight before the first store waits on one
  for (int mb = 0; mb < tt.n; mb += 4) {
    double2 v[4][2];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int m = min(mb + u, tt.n - 1);
      v[u][0] = *reinterpret_cast<const double2 *>(tt.src[m] + q0);
      v[u][1] = *reinterpret_cast<const double2 *>(tt.src[m] + q0 + 2);
    }
#pragma unroll
    for (int u = 0; u < 4; u++) {
      const int m = mb + u;
      if (m < tt.n) {
        double o[4] = {v[u][0].x, v[u][0].y, v[u][1].x, v[u][1].y};
        const double cu = tt.cu[m];
        if (cu > 0.0) {
#pragma unroll
          for (int n = 0; n < 4; n++) if (inr[n]) o[n] = underflow(o[n], cu);
        }
        *reinterpret_cast<double2 *>(tt.dst[m] + q0) = make_double2(o[0], o[1]);
        *reinterpret_cast<double2 *>(tt.dst[m] + q0 + 2) = make_double2(o[2], o[3]);
      }
    }
  }
}

#ifndef XL2PF
#define XL2PF 6        // tracers ahead of the computed one whose lines are pulled into L2 (0: off); measured 3: 21.7, 5-8: 21.1 ms
#endif
#ifndef XMINB
#define XMINB 4
#endif
// DIAG: the tracers of this launch have ad_x and / or advection_xy associated (dg, same order as tt): the face fluxes
// sweep4 leaves in registers are accumulated where the reference does (:665-667, :671-676).  The plain instantiation
// (every other tracer) is unchanged by it.
template <int SCHEME, bool DIAG>
__global__ void __launch_bounds__(128, XMINB)
k_x_fast(Geom g, Range rg, TrTable tt, DiagTable dg, const double *__restrict__ IareaT, double Idt,
         const double *__restrict__ hhv, const double *__restrict__ hcv,
         const double *__restrict__ gqv, const double *__restrict__ hlstv, const double *__restrict__ ihv,
         const double *__restrict__ mmv, const int *__restrict__ entry, const int *__restrict__ domore_k, int wbase,
         int wj_lo, int wj_hi,        // rows (Fortran j) this launch owns; the others belong to another launch of the sweep
         int edge_store) {            // tiles aligned on the sweep's own range: the outermost halo lanes copy the columns beside it
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int rj = wj_lo + g.roff + blockIdx.y * 4 + (threadIdx.x >> 5), k0 = blockIdx.z;
  const int j = rj - g.roff;
  if (j > wj_hi) return;                                     // warp-uniform
  const int c0 = wbase + XT_USE * blockIdx.x + 4 * lane;       // first of this lane's 4 columns
  const long long q0 = (long long)c0 + (long long)g.P * rj + g.plane * k0;
  const bool lev = domore_k[k0] > 0;
  const bool in_j = (j >= rg.js && j <= rg.je);
  const bool active = lev && in_j && entry[k0 * g.R + rj];
  // lanes 1..30 own their cells; with edge_store lane 0 of the first tile and lane 31 of the last one also store theirs:
  // those columns lie outside the sweep's range (never updated, output == input), so no further tile is spent on them
  const bool edge = edge_store && ((lane == 0 && blockIdx.x == 0) || (lane == 31 && blockIdx.x == gridDim.x - 1));
  const bool st = ((lane >= 1 && lane <= 30) || edge) && (c0 + 3 >= g.isd + g.coff) && (c0 <= g.ied + g.coff);
  bool inr[4];
#pragma unroll
  for (int n = 0; n < 4; n++) { const int i = c0 + n - g.coff; inr[n] = lev && in_j && i >= rg.is && i <= rg.ie; }

  if (!active) { x_copy_through(tt, q0, st, inr); return; }

  // tracer-independent quantities of this lane's 4 cells and their 5 faces
  double hh[5], hc[5], gq[5], hl[4], ih[4], mm[4];     // faces: [0] = west face of cell 0, [n+1] = east face of cell n
  {
    const double2 a0 = *reinterpret_cast<const double2 *>(hhv + q0), a1 = *reinterpret_cast<const double2 *>(hhv + q0 + 2);
    const double2 b0 = *reinterpret_cast<const double2 *>(hcv + q0), b1 = *reinterpret_cast<const double2 *>(hcv + q0 + 2);
    const double2 c0v = *reinterpret_cast<const double2 *>(gqv + q0), c1v = *reinterpret_cast<const double2 *>(gqv + q0 + 2);
    const double2 h0 = *reinterpret_cast<const double2 *>(hlstv + q0), h1 = *reinterpret_cast<const double2 *>(hlstv + q0 + 2);
    const double2 i0 = *reinterpret_cast<const double2 *>(ihv + q0), i1 = *reinterpret_cast<const double2 *>(ihv + q0 + 2);
    const long long q2 = (long long)c0 + (long long)g.P * rj;
    const double2 m0 = *reinterpret_cast<const double2 *>(mmv + q2), m1 = *reinterpret_cast<const double2 *>(mmv + q2 + 2);
    hh[1] = a0.x; hh[2] = a0.y; hh[3] = a1.x; hh[4] = a1.y;
    hc[1] = b0.x; hc[2] = b0.y; hc[3] = b1.x; hc[4] = b1.y;
    gq[1] = c0v.x; gq[2] = c0v.y; gq[3] = c1v.x; gq[4] = c1v.y;
    hl[0] = h0.x; hl[1] = h0.y; hl[2] = h1.x; hl[3] = h1.y;
    ih[0] = i0.x; ih[1] = i0.y; ih[2] = i1.x; ih[3] = i1.y;
    mm[0] = m0.x; mm[1] = m0.y; mm[2] = m1.x; mm[3] = m1.y;
    hh[0] = __shfl_up_sync(FULL, hh[4], 1);
    hc[0] = __shfl_up_sync(FULL, hc[4], 1);
    gq[0] = __shfl_up_sync(FULL, gq[4], 1);
  }
  // per-lane flags, tracer-independent, packed so that they cost one register instead of 13 predicates:
  // bit n (0..4): face n is fed by its low-side cell (:519 uhh >= 0); bit 5+n: cell n is updated (:657 do_i,
  // Ihnew > 0); bit 9+n: cell n is inside the sweep's own range (underflow test, :683-687)
  unsigned flg = 0;
#pragma unroll
  for (int n = 0; n < 5; n++) flg |= (hh[n] >= 0.0) ? (1u << n) : 0u;
#pragma unroll
  for (int n = 0; n < 4; n++) flg |= ((ih[n] > 0.0) ? (32u << n) : 0u) | (inr[n] ? (512u << n) : 0u);
  // no stored cell of this warp's 128 columns is updated (land, or no transport left on any of their faces):
  // every output equals its input whatever the fluxes are, so the tile is copied through
  // (DIAG: the east face of lane 30's last cell also counts when lane 31's first cell -- a stored cell of the NEXT tile --
  // is updated: ad_x accumulates where do_i(i) .or. do_i(i+1), :665)
  if (!__any_sync(FULL, (st && (flg & 0x1e0u) != 0u) || (DIAG && lane == 31 && (flg & 0x20u) != 0u))) { x_copy_through(tt, q0, st, inr); return; }
  const unsigned de_flg = DIAG ? __shfl_down_sync(FULL, flg, 1) : 0u;
  const bool st_ok = st && !(edge && (flg & 0x1e0u) != 0u);      // (an edge lane never holds an updated cell; if it did, its value would need a neighbour it lacks)
  asm volatile("" : "+r"(flg));      // opaque: keeps the compiler from re-deriving the bits with FP64 compares per tracer
  // signed flux factors (face_flux_sgn): PPM wants -hc, -gq on faces fed from the low side, PLM +gq there
  double hcs[5], gqs[5];
#pragma unroll
  for (int n = 0; n < 5; n++) {
    const bool lo = (flg >> n) & 1u;
    hcs[n] = neg_if(hc[n], lo);
    gqs[n] = neg_if(gq[n], SCHEME == SCHEME_PLM ? !lo : lo);
  }

  double ia[4] = {0, 0, 0, 0};
  if (DIAG) {
    const long long q2 = (long long)c0 + (long long)g.P * rj;
    const double2 a0 = *reinterpret_cast<const double2 *>(IareaT + q2), a1 = *reinterpret_cast<const double2 *>(IareaT + q2 + 2);
    ia[0] = a0.x; ia[1] = a0.y; ia[2] = a1.x; ia[3] = a1.y;
  }
  // two tracers in flight ahead of the one being computed: one iteration of the loop is shorter than a loaded
  // HBM round trip (ncu: 17 % of the samples sat on the first use of a one-deep prefetch)
  double2 n0 = *reinterpret_cast<const double2 *>(tt.src[0] + q0);
  double2 n1 = *reinterpret_cast<const double2 *>(tt.src[0] + q0 + 2);
  const int m1 = (tt.n > 1) ? 1 : 0;
  double2 p0 = *reinterpret_cast<const double2 *>(tt.src[m1] + q0);
  double2 p1 = *reinterpret_cast<const double2 *>(tt.src[m1] + q0 + 2);
  for (int m = 0; m < tt.n; m++) {
    double T[6];      // T[1..4] own cells (sweep4 fetches the neighbours T[0], T[5] from the adjacent lanes)
    T[1] = n0.x; T[2] = n0.y; T[3] = n1.x; T[4] = n1.y;
    n0 = p0; n1 = p1;
    if (m + 2 < tt.n) {       // prefetch the tracer after the next
      p0 = *reinterpret_cast<const double2 *>(tt.src[m + 2] + q0);
      p1 = *reinterpret_cast<const double2 *>(tt.src[m + 2] + q0 + 2);
    }
#if XL2PF > 0
    if (m + XL2PF < tt.n)     // and pull a later tracer's lines into L2: costs no register, shortens the round trip above
      asm volatile("prefetch.global.L2 [%0];" ::"l"(tt.src[m + XL2PF] + q0));
#endif
    double o[4];
    if (DIAG) {
      double fl[5];
      sweep4<SCHEME, false>(T, hh, hcs, gqs, hl, ih, mm, flg, o, fl);
      if (st_ok) {
        // do_i of this lane's cells (bits 5..8) and of the east neighbour's first cell (de_flg, fetched before the loop)
        const unsigned dn = (flg >> 5) & 0xfu;
        const unsigned de = (de_flg >> 5) & 1u;
        const unsigned dnx = (dn >> 1) | ((lane == 31 ? 0u : de) << 3);      // do_i of cell n+1
        double *__restrict__ ad = dg.ad[m];
        double *__restrict__ axy = dg.axy[m];
#pragma unroll
        for (int n = 0; n < 4; n++) {
          // face I = east face of cell n, stored at the cell's own index: do_i(i) .or. do_i(i+1)  (:665-667)
          if (ad && (((dn | dnx) >> n) & 1u)) ad[q0 + n] = ad[q0 + n] + fl[n + 1] * Idt;
          if (axy && ((dn >> n) & 1u)) axy[q0 + n] = axy[q0 + n] - (fl[n + 1] - fl[n]) * Idt * ia[n];      // :671-676
        }
      }
    } else {
      sweep4<SCHEME, false>(T, hh, hcs, gqs, hl, ih, mm, flg, o);
    }
    const double cu = tt.cu[m];
    if (cu > 0.0) {      // warp-uniform: most tracers carry no underflow threshold
#pragma unroll
      for (int n = 0; n < 4; n++) if ((flg >> (9 + n)) & 1u) o[n] = underflow(o[n], cu);
    }
    if (st_ok) {
      *reinterpret_cast<double2 *>(tt.dst[m] + q0) = make_double2(o[0], o[1]);
      *reinterpret_cast<double2 *>(tt.dst[m] + q0 + 2) = make_double2(o[2], o[3]);
    }
  }
}

int x_fast_tiles(const Geom &g, int wbase) { return (g.ied + g.coff - (wbase + 4)) / XT_USE + 1; }

void launch_x_fast(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *hc, const double *gq, const double *hlst, const double *ihnew, const double *mm,
                   const int *entry, const int *domore_k, int wbase, int wj_lo, int wj_hi,
                   const DiagTable *dg, const double *IareaT, double Idt) {
  wj_lo = std::max(wj_lo, g.jsd); wj_hi = std::min(wj_hi, g.jed);
  if (wj_lo > wj_hi) return;
  int ntx = x_fast_tiles(g, wbase), edge = 0;
  {
    // Tiles aligned on the sweep's own columns [rg.is, rg.ie] instead of the data domain: with halo 4 and one stencil
    // application per pass the sweep computes isc..iec only, the 4 + 4 halo columns are copied by the two outermost
    // halo lanes, and OM4_025's 1440 columns take 12 tiles instead of 13 (the 13th computed 4 cells).
    const int cs = rg.is + g.coff, ce = rg.ie + g.coff;
    const int wb2 = (cs & ~3) - 4;
    const int nt2 = (ce >= cs) ? (ce - (wb2 + 4)) / XT_USE + 1 : ntx;
    if (ce >= cs && wb2 >= 0 && nt2 < ntx && g.isd + g.coff >= wb2 && g.ied + g.coff <= wb2 + XT_USE * nt2 + 7 &&
        wb2 + XT_USE * (nt2 - 1) + 128 <= g.P) { wbase = wb2; ntx = nt2; edge = 1; }
  }
  dim3 grid(ntx, (wj_hi - wj_lo + 1 + 3) / 4, g.nk);
  static const DiagTable none = {};
#define XL(S)                                                                                                          \
  do {                                                                                                                 \
    if (dg) k_x_fast<S, true><<<grid, 128, 0, s>>>(g, rg, tt, *dg, IareaT, Idt, hh, hc, gq, hlst, ihnew, mm, entry, domore_k, wbase, wj_lo, wj_hi, edge); \
    else k_x_fast<S, false><<<grid, 128, 0, s>>>(g, rg, tt, none, nullptr, 0.0, hh, hc, gq, hlst, ihnew, mm, entry, domore_k, wbase, wj_lo, wj_hi, edge); \
  } while (0)
  if (scheme == SCHEME_PLM) XL(SCHEME_PLM);
  else if (scheme == SCHEME_PPM_CW) XL(SCHEME_PPM_CW);
  else XL(SCHEME_PPM_H3);
#undef XL
}
