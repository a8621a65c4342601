// TEST INFRASTRUCTURE ONLY -- C entry point around the code oracle/f2c/f2cpp.py generates from the unmodified
// /root/reference/src/tracer/MOM_tracer_advect.F90 (advect_tracer, advect_x, advect_y), with the argument list of
// ora_advect_tracer_obc (oracle/advect_ref.h) so that tests can run the hand-written restatement and the mechanical
// translation of the reference side by side on the same inputs.  Built into oracle/_ref/libref_f2c*.so by
// oracle/f2c/Makefile; the group pass (FMS data movement, not part of the reference file) is the oracle's emulation,
// exactly the role SURVEY.md appendix B gives the MOM_domains stub.
#include "f2c_runtime.h"
#include "../advect_ref.h"
#include "obc_from_ora.h"

thread_local f2c_host *f2c_current_host = nullptr;

#include "advect_f2c.inc"      // generated (oracle/_ref/)

namespace {
struct HostCtx {
  ora_pass_fn pass; ora_sum_fn sum; void *ctx;
  ora_tracer *Tr; int ntr;
  ora_rank rank; const ora_domain *dom;
};
void host_pass(void *c, group_pass_type &g) {
  HostCtx *h = (HostCtx *)c;
  if (h->pass) { h->pass(h->ctx, g.u, g.v, g.scalars.at(0), h->Tr, h->ntr); return; }
  // built-in single-rank pass: the oracle's full-width CGRID_NE emulation
  double *u[1] = {g.u}, *v[1] = {g.v}, *hp[1] = {g.scalars.at(0)};
  h->rank.Tr = h->Tr;
  ora_group_pass(1, &h->rank, h->dom, u, v, hp, h->ntr);
}
void host_sum(void *c, int *v, int n) { HostCtx *h = (HostCtx *)c; if (h->sum) h->sum(h->ctx, v, n); }
}  // namespace

extern "C" int ref_f2c_logical_fill(void) { return F2C_LOGICAL_FILL; }
extern "C" int ref_f2c_minmax_later_wins(void) { return F2C_MINMAX_LATER_WINS; }

extern "C" int ref_advect_tracer_obc(const ora_grid *G, const ora_cs *CS, ora_tracer *Tr, int ntr, const double *h_end,
                                     const double *uhtr, const double *vhtr, double dt, int x_first_in, double *vol_prev,
                                     int max_iter_in, int update_vol_prev, double *uhr_out, double *vhr_out,
                                     const ora_domain *dom, ora_pass_fn pass, ora_sum_fn sum, void *ctx, int *n_pass_out,
                                     const ora_obc *OBC) {
  ocean_grid_type g;
  g.isc = G->isc; g.iec = G->iec; g.jsc = G->jsc; g.jec = G->jec;
  g.isd = G->isd; g.ied = G->ied; g.jsd = G->jsd; g.jed = G->jed;
  g.isdb = G->IsdB; g.iedb = G->IedB; g.jsdb = G->JsdB; g.jedb = G->JedB;
  g.ke = G->nk; g.first_direction = G->first_direction;
  mom_domain_type domain; g.domain = &domain;
  const Bnd hi_ = {G->isd, G->ied}, hj_ = {G->jsd, G->jed}, ui_ = {G->IsdB, G->IedB}, vj_ = {G->JsdB, G->JedB}, kk = {1, G->nk};
  g.areat = Arr<double, 2>::view(const_cast<double *>(G->areaT), {hi_, hj_});
  g.iareat = Arr<double, 2>::view(const_cast<double *>(G->IareaT), {hi_, hj_});
  g.mask2dt = Arr<double, 2>::view(const_cast<double *>(G->mask2dT), {hi_, hj_});
  g.mask2dcu = Arr<double, 2>::view(const_cast<double *>(G->mask2dCu), {ui_, hj_});
  g.mask2dcv = Arr<double, 2>::view(const_cast<double *>(G->mask2dCv), {hi_, vj_});
  verticalgrid_type gv; gv.ke = G->nk; gv.angstrom_h = G->Angstrom_H; gv.h_subroundoff = G->H_subroundoff;
  unit_scale_type us;
  tracer_advect_cs cs; cs.dt = CS->dt; cs.useppm = CS->usePPM != 0; cs.usehuynh = CS->useHuynh != 0;
  cs.usehuynhstencilbug = CS->useHuynhStencilBug != 0;
  tracer_registry_type reg; reg.ntr = ntr;
  reg.tr = Arr<tracer_type, 1>::automatic({{1, ntr > 0 ? ntr : 1}});
  for (int m = 1; m <= ntr; m++) {
    tracer_type &t = reg.tr(m); const ora_tracer &s = Tr[m - 1];
    t.t = Arr<double, 3>::view(s.t, {hi_, hj_, kk});
    t.ad_x = Arr<double, 3>::view(s.ad_x, {ui_, hj_, kk});
    t.ad_y = Arr<double, 3>::view(s.ad_y, {hi_, vj_, kk});
    t.advection_xy = Arr<double, 3>::view(s.advection_xy, {hi_, hj_, kk});
    t.ad2d_x = Arr<double, 2>::view(s.ad2d_x, {ui_, hj_});
    t.ad2d_y = Arr<double, 2>::view(s.ad2d_y, {hi_, vj_});
    t.conc_underflow = s.conc_underflow;
  }
  // open boundaries
  ObcHolder obc_holder;
  ocean_obc_type *obcp = obc_from_ora(OBC, obc_holder, G->nk);
  HostCtx hc; hc.pass = pass; hc.sum = sum; hc.ctx = ctx; hc.Tr = Tr; hc.ntr = ntr; hc.dom = dom;
  memset(&hc.rank, 0, sizeof hc.rank); hc.rank.G = *G;
  hc.rank.ioff = -(G->isc - 1); hc.rank.joff = -(G->jsc - 1);      // single block: local isc is global 1
  if (!pass && !dom) return 1;
  f2c_host host; host.pass = host_pass; host.sum = host_sum; host.ctx = &hc;
  f2c_current_host = &host;
  int rc = 0;
  try {
    advect_tracer(Arr<double, 3>::view(const_cast<double *>(h_end), {hi_, hj_, kk}),
                  Arr<double, 3>::view(const_cast<double *>(uhtr), {ui_, hj_, kk}),
                  Arr<double, 3>::view(const_cast<double *>(vhtr), {hi_, vj_, kk}), obcp, dt, g, gv, us, &cs, &reg,
                  x_first_in < 0 ? Opt<lbool>() : Opt<lbool>((lbool)(x_first_in != 0)),
                  Arr<double, 3>::view(vol_prev, {hi_, hj_, kk}),
                  max_iter_in < 0 ? Opt<int>() : Opt<int>(max_iter_in),
                  update_vol_prev < 0 ? Opt<lbool>() : Opt<lbool>((lbool)(update_vol_prev != 0)),
                  Arr<double, 3>::view(uhr_out, {ui_, hj_, kk}), Arr<double, 3>::view(vhr_out, {hi_, vj_, kk}));
  } catch (const std::string &) {
    rc = 1;
  }
  f2c_current_host = nullptr;
  if (n_pass_out) *n_pass_out = host.n_pass;
  return rc;
}
