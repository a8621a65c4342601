// Halo-exchange planning for the multi-GPU group pass: the device-resident
// replacement of do_group_pass(CS%pass_uhr_vhr_t_hprev, G%Domain)
// (/root/reference/src/tracer/MOM_tracer_advect.F90:140-144,206 through
// config_src/infra/FMS2/MOM_domain_infra.F90:983-1158, i.e. FMS
// mpp_do_group_update): full-width halos, all four sides and the corners,
// scalars hprev and every tracer, and the C-grid vector pair (uhr, vhr) whose
// components change sign across a tripolar north fold.  Cyclic wrap and the fold
// are index maps applied when the lists are built; points without a source
// (closed boundaries) are left untouched, as FMS does.
//
// Every rank enumerates the halo points of every receiver in the same order, so
// sender and receiver agree on the message layout without exchanging lists.
#include "../../include/mom6adv.h"
#include "internal.cuh"
#include "kernels.h"

#include <vector>

namespace {

struct RankGeo { int isc, iec, jsc, jec, isd, ied, jsd, jed, IsdB, JsdB, ioff, joff; };

// local index space of the block (pi,pj): isd = jsd = 1 (MOM_domain_infra.F90:1815-1821)
RankGeo rank_geo(const ExchangeSpec &x, int pi, int pj) {
  RankGeo r;
  const int ni = x.iend[pi] - x.ibeg[pi] + 1, nj = x.jend[pj] - x.jbeg[pj] + 1;
  r.isd = 1; r.jsd = 1;
  r.isc = 1 + x.halo_i; r.jsc = 1 + x.halo_j;
  r.iec = r.isc + ni - 1; r.jec = r.jsc + nj - 1;
  r.ied = r.iec + x.halo_i; r.jed = r.jec + x.halo_j;
  r.IsdB = x.symmetric ? r.isd - 1 : r.isd;
  r.JsdB = x.symmetric ? r.jsd - 1 : r.jsd;
  r.ioff = x.ibeg[pi] - r.isc;      // global = local + off
  r.joff = x.jbeg[pj] - r.jsc;
  return r;
}

struct Src { int rank, i, j, sgn; };

// Owner of global point (ig,jg) of stagger pos (0 h, 1 u, 2 v); false at a closed boundary.
bool locate(const ExchangeSpec &x, const std::vector<RankGeo> &geo, int pos, int ig, int jg, Src *s) {
  const int nig = x.nig, njg = x.njg;
  int sgn = 1;
  if (x.tripolar_n && jg > njg) {      // fold line = v-face row njg
    if (pos == 0) { ig = nig + 1 - ig; jg = 2 * njg + 1 - jg; }
    else if (pos == 1) { ig = nig - ig; jg = 2 * njg + 1 - jg; sgn = -1; }
    else { ig = nig + 1 - ig; jg = 2 * njg - jg; sgn = -1; }
  }
  const int ilo = (pos == 1) ? 0 : 1, jlo = (pos == 2) ? 0 : 1;
  if (x.cyclic_x) { while (ig < 1) ig += nig; while (ig > nig) ig -= nig; }
  else if (ig < ilo || ig > nig) return false;
  if (x.cyclic_y) { while (jg < 1) jg += njg; while (jg > njg) jg -= njg; }
  else if (jg < jlo || jg > njg) return false;
  for (int r = 0; r < (int)geo.size(); r++) {
    const RankGeo &G = geo[r];
    const int i = ig - G.ioff, j = jg - G.joff;
    int i0 = G.isc, j0 = G.jsc;
    if (pos == 1 && ig == 0) i0 = G.isc - 1;     // western domain-boundary face
    if (pos == 2 && jg == 0) j0 = G.jsc - 1;     // southern domain-boundary face
    if (i >= i0 && i <= G.iec && j >= j0 && j <= G.jec) { s->rank = r; s->i = i; s->j = j; s->sgn = sgn; return true; }
  }
  return false;
}

}  // namespace

// Build, for rank `me`, the self-sourced halo list and the per-peer send / receive lists.
// Padded plane index of local (i,j): (i + g.coff) + g.P * (j + g.roff).
void build_halo_exchange(const Geom &g, const ExchangeSpec &x, int me, HaloBuild local[3], std::vector<PeerBuild> &peers) {
  const int nr = x.npi * x.npj;
  std::vector<RankGeo> geo(nr);
  for (int pj = 0; pj < x.npj; pj++) for (int pi = 0; pi < x.npi; pi++) geo[pi + x.npi * pj] = rank_geo(x, pi, pj);
  std::vector<int> slot(nr, -1);
  auto peer = [&](int r) -> PeerBuild & {
    if (slot[r] < 0) { slot[r] = (int)peers.size(); peers.emplace_back(); peers.back().rank = r; }
    return peers[slot[r]];
  };
  auto pidx = [&](int i, int j) { return (i + g.coff) + g.P * (j + g.roff); };
  // padded layouts of the other ranks (make_geom is a pure function of the index bounds)
  std::vector<Geom> pg(nr);
  for (int r = 0; r < nr; r++) {
    mom6adv_grid gr = {};
    gr.isc = geo[r].isc; gr.iec = geo[r].iec; gr.jsc = geo[r].jsc; gr.jec = geo[r].jec;
    gr.isd = geo[r].isd; gr.ied = geo[r].ied; gr.jsd = geo[r].jsd; gr.jed = geo[r].jed;
    gr.IsdB = geo[r].IsdB; gr.IedB = geo[r].ied; gr.JsdB = geo[r].JsdB; gr.JedB = geo[r].jed; gr.nk = g.nk;
    int wb;
    make_geom(&gr, pg[r], wb);
  }
  for (int d = 0; d < nr; d++) {            // receiver
    const RankGeo &G = geo[d];
    for (int pos = 0; pos < 3; pos++) {
      const int ilo = (pos == 1) ? G.IsdB : G.isd, jlo = (pos == 2) ? G.JsdB : G.jsd;
      int ic0 = G.isc, jc0 = G.jsc;
      if (x.symmetric) { if (pos == 1) ic0 = G.isc - 1; if (pos == 2) jc0 = G.jsc - 1; }
      for (int j = jlo; j <= G.jed; j++) for (int i = ilo; i <= G.ied; i++) {
        if (i >= ic0 && i <= G.iec && j >= jc0 && j <= G.jec) continue;     // not a halo point
        if (!halo_point_used(pos, i, j, G.isc, G.iec, G.jsc, G.jec, x.depth)) continue;
        Src s;
        if (!locate(x, geo, pos, i + G.ioff, j + G.joff, &s)) continue;
        if (d == me && s.rank == me) {
          local[pos].dst.push_back(pidx(i, j)); local[pos].src.push_back(pidx(s.i, s.j));
          local[pos].sgn.push_back((signed char)s.sgn);
        } else if (d == me) {
          PeerBuild &p = peer(s.rank);
          p.recv_dst[pos].push_back(pidx(i, j));
          const Geom &q = pg[s.rank];
          p.recv_src[pos].push_back((s.i + q.coff) + q.P * (s.j + q.roff));
          p.recv_sgn[pos].push_back((signed char)s.sgn);
          p.peer_plane = q.plane; p.peer_P = q.P; p.peer_R = q.R;
        } else if (s.rank == me) {
          PeerBuild &p = peer(d);
          p.send_src[pos].push_back(pidx(s.i, s.j)); p.send_sgn[pos].push_back((signed char)s.sgn);
        }
      }
    }
  }
}
