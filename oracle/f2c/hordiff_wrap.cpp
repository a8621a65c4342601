// TEST INFRASTRUCTURE ONLY -- the along-layer diffusion loop of tracer_hordiff,
// /root/reference/src/tracer/MOM_tracer_hor_diff.F90:539-607, UNMODIFIED, translated statement by statement by
// oracle/f2c/f2cpp.py --range (the subroutine around it -- MEKE, VarMix, neutral and boundary diffusion, epipycnal
// mixing -- is not translated: this wrapper declares the loop's variables with the reference's declared shapes,
// :150-180, and includes the generated statements).  Same argument list as ora_tracer_hordiff (oracle/advect_ref.h).
#include "f2c_runtime.h"
#include "../advect_ref.h"

extern thread_local f2c_host *f2c_current_host;

namespace {
struct HdCtx { ora_pass_fn pass; void *ctx; ora_tracer *Tr; int ntr; ora_rank rank; const ora_domain *dom;
               std::vector<double> su, sv, sh; };
void hd_pass(void *c, group_pass_type &g) {
  HdCtx *h = (HdCtx *)c;
  (void)g;      // CS%pass_t holds the tracers (create_group_pass per tracer, MOM_tracer_hor_diff.F90:399-403)
  if (h->pass) { h->pass(h->ctx, h->su.data(), h->sv.data(), h->sh.data(), h->Tr, h->ntr); return; }
  double *u[1] = {h->su.data()}, *v[1] = {h->sv.data()}, *hp[1] = {h->sh.data()};
  h->rank.Tr = h->Tr;
  ora_group_pass(1, &h->rank, h->dom, u, v, hp, h->ntr);
}
}  // namespace

static int hordiff_ref_impl(const ora_grid *G, ora_tracer *Tr, int ntr_in, const double *h_in, const double *khdt_x_in,
                            const double *khdt_y_in, int num_itts, double h_neglect, int Diffuse_ML_interior,
                            int nkml, int nk_rho_varies, double ML_KhTr_scale, double dt, const ora_domain *dom,
                            ora_pass_fn pass, void *ctx, double *const *df_x, double *const *df_y,
                            double *const *df2d_x, double *const *df2d_y) {
  if (!pass && !dom) return 1;
  ocean_grid_type g;
  g.isc = G->isc; g.iec = G->iec; g.jsc = G->jsc; g.jec = G->jec;
  g.isd = G->isd; g.ied = G->ied; g.jsd = G->jsd; g.jed = G->jed;
  g.isdb = G->IsdB; g.iedb = G->IedB; g.jsdb = G->JsdB; g.jedb = G->JedB; g.ke = G->nk;
  // MOM_hor_index.F90:87-96: symmetric memory has IscB = isc-1, else IscB = isc
  g.iscb = (G->IsdB == G->isd - 1) ? G->isc - 1 : G->isc; g.iecb = G->iec;
  g.jscb = (G->JsdB == G->jsd - 1) ? G->jsc - 1 : G->jsc; g.jecb = G->jec;
  mom_domain_type domain; g.domain = &domain;
  const Bnd hi_ = {G->isd, G->ied}, hj_ = {G->jsd, G->jed}, ui_ = {G->IsdB, G->IedB}, vj_ = {G->JsdB, G->JedB}, kk = {1, G->nk};
  g.iareat = Arr<double, 2>::view(const_cast<double *>(G->IareaT), {hi_, hj_});
  verticalgrid_type gv; gv.ke = G->nk; gv.h_subroundoff = G->H_subroundoff; gv.nkml = nkml; gv.nk_rho_varies = nk_rho_varies;
  tracer_hor_diff_cs cs_; tracer_hor_diff_cs *cs = &cs_;
  cs->diffuse_ml_interior = Diffuse_ML_interior != 0; cs->ml_khtr_scale = ML_KhTr_scale;
  tracer_registry_type reg_; tracer_registry_type *reg = &reg_;
  reg->ntr = ntr_in;
  reg->tr = Arr<tracer_type, 1>::automatic({{1, ntr_in > 0 ? ntr_in : 1}});
  for (int m = 1; m <= ntr_in; m++) {
    reg->tr(m).t = Arr<double, 3>::view(Tr[m - 1].t, {hi_, hj_, kk});
    reg->tr(m).conc_underflow = Tr[m - 1].conc_underflow;
    // the diffusive-flux diagnostics, associated where the caller hands an array (MOM_tracer_types.F90 shapes)
    if (df_x && df_x[m - 1]) reg->tr(m).df_x = Arr<double, 3>::view(df_x[m - 1], {ui_, hj_, kk});
    if (df_y && df_y[m - 1]) reg->tr(m).df_y = Arr<double, 3>::view(df_y[m - 1], {hi_, vj_, kk});
    if (df2d_x && df2d_x[m - 1]) reg->tr(m).df2d_x = Arr<double, 2>::view(df2d_x[m - 1], {ui_, hj_});
    if (df2d_y && df2d_y[m - 1]) reg->tr(m).df2d_y = Arr<double, 2>::view(df2d_y[m - 1], {hi_, vj_});
  }
  // the subroutine's own variables, with the shapes it declares them with (:150-180)
  Arr<double, 3> h = Arr<double, 3>::view(const_cast<double *>(h_in), {hi_, hj_, kk});
  Arr<double, 2> khdt_x = Arr<double, 2>::view(const_cast<double *>(khdt_x_in), {ui_, hj_});
  Arr<double, 2> khdt_y = Arr<double, 2>::view(const_cast<double *>(khdt_y_in), {hi_, vj_});
  Arr<double, 2> ihdxdy = Arr<double, 2>::automatic({hi_, hj_}), dtr = Arr<double, 2>::automatic({hi_, hj_});
  Arr<double, 3> coef_x = Arr<double, 3>::automatic({ui_, hj_, {1, G->nk + 1}});
  Arr<double, 3> coef_y = Arr<double, 3>::automatic({hi_, vj_, {1, G->nk + 1}});
  int i = 0, j = 0, k = 0, m = 0, itt = 0;
  const int is = G->isc, ie = G->iec, js = G->jsc, je = G->jec, nz = G->nk, ntr = ntr_in;      // :189-190
  double scale = f_uninit<double>();
  const double i_numitts = 1.0 / ((double)num_itts);      // :380 / :384 / :386
  const double idt = 1.0 / dt;                              // :204
  (void)idt; (void)m; (void)i; (void)j; (void)k; (void)itt;

  HdCtx hc; hc.pass = pass; hc.ctx = ctx; hc.Tr = Tr; hc.ntr = ntr_in; hc.dom = dom;
  memset(&hc.rank, 0, sizeof hc.rank); hc.rank.G = *G; hc.rank.ioff = -(G->isc - 1); hc.rank.joff = -(G->jsc - 1);
  hc.su.assign((size_t)(G->IedB - G->IsdB + 1) * (G->jed - G->jsd + 1) * G->nk, 0.0);
  hc.sv.assign((size_t)(G->ied - G->isd + 1) * (G->JedB - G->JsdB + 1) * G->nk, 0.0);
  hc.sh.assign((size_t)(G->ied - G->isd + 1) * (G->jed - G->jsd + 1) * G->nk, 0.0);
  f2c_host host; host.pass = hd_pass; host.ctx = &hc;
  f2c_current_host = &host;
  int rc = 0;
  try {
#include "hordiff_f2c.inc"      // generated: MOM_tracer_hor_diff.F90:539-607
  } catch (const std::string &) { rc = 1; }
  f2c_current_host = nullptr;
  return rc;
}

extern "C" int ref_tracer_hordiff(const ora_grid *G, ora_tracer *Tr, int ntr_in, const double *h_in, const double *khdt_x_in,
                                  const double *khdt_y_in, int num_itts, double h_neglect, int Diffuse_ML_interior,
                                  int nkml, int nk_rho_varies, double ML_KhTr_scale, double dt, const ora_domain *dom,
                                  ora_pass_fn pass, void *ctx) {
  return hordiff_ref_impl(G, Tr, ntr_in, h_in, khdt_x_in, khdt_y_in, num_itts, h_neglect, Diffuse_ML_interior, nkml,
                          nk_rho_varies, ML_KhTr_scale, dt, dom, pass, ctx, nullptr, nullptr, nullptr, nullptr);
}

// with the df_x / df_y / df2d_x / df2d_y diagnostics of :576-591 (same conventions as ora_tracer_hordiff_df)
extern "C" int ref_tracer_hordiff_df(const ora_grid *G, ora_tracer *Tr, int ntr_in, const double *h_in, const double *khdt_x_in,
                                     const double *khdt_y_in, int num_itts, double h_neglect, int Diffuse_ML_interior,
                                     int nkml, int nk_rho_varies, double ML_KhTr_scale, double dt, const ora_domain *dom,
                                     ora_pass_fn pass, void *ctx, double *const *df_x, double *const *df_y,
                                     double *const *df2d_x, double *const *df2d_y) {
  return hordiff_ref_impl(G, Tr, ntr_in, h_in, khdt_x_in, khdt_y_in, num_itts, h_neglect, Diffuse_ML_interior, nkml,
                          nk_rho_varies, ML_KhTr_scale, dt, dom, pass, ctx, df_x, df_y, df2d_x, df2d_y);
}

// num_itts / I_numitts from the diffusive CFL, MOM_tracer_hor_diff.F90:368-387 (translated; one PE: max_across_PEs is
// the identity; the call-tree, clock and post_data calls do nothing)
namespace {
template <class... A> void calltree_waypoint(A &&...) {}
template <class... A> void cpu_clock_begin(A &&...) {}
template <class... A> void cpu_clock_end(A &&...) {}
template <class... A> void post_data(A &&...) {}
template <class T> T kw_mask(T v) { return v; }
inline void max_across_pes(double &) {}
}  // namespace

extern "C" int ref_hordiff_num_itts(const ora_grid *G, const double *khdt_x_in, const double *khdt_y_in,
                                    int check_diffusive_CFL, double max_diff_CFL, int *num_itts_out, double *I_numitts_out) {
  ocean_grid_type g;
  g.isc = G->isc; g.iec = G->iec; g.jsc = G->jsc; g.jec = G->jec;
  g.isd = G->isd; g.ied = G->ied; g.jsd = G->jsd; g.jed = G->jed;
  const Bnd hi_ = {G->isd, G->ied}, hj_ = {G->jsd, G->jed}, ui_ = {G->IsdB, G->IedB}, vj_ = {G->JsdB, G->JedB};
  g.iareat = Arr<double, 2>::view(const_cast<double *>(G->IareaT), {hi_, hj_});
  g.mask2dt = Arr<double, 2>::view(const_cast<double *>(G->mask2dT), {hi_, hj_});
  tracer_hor_diff_cs cs_; tracer_hor_diff_cs *cs = &cs_;
  cs->check_diffusive_cfl = check_diffusive_CFL != 0; cs->max_diff_cfl = max_diff_CFL;
  Arr<double, 2> khdt_x = Arr<double, 2>::view(const_cast<double *>(khdt_x_in), {ui_, hj_});
  Arr<double, 2> khdt_y = Arr<double, 2>::view(const_cast<double *>(khdt_y_in), {hi_, vj_});
  Arr<double, 2> cfl = Arr<double, 2>::automatic({hi_, hj_});      // CFL(SZI_(G),SZJ_(G)), :140
  const int is = G->isc, ie = G->iec, js = G->jsc, je = G->jec;
  int i = 0, j = 0, num_itts = 0, id_clock_sync = 0;
  double max_cfl = f_uninit<double>(), i_numitts = f_uninit<double>();
  (void)i; (void)j; (void)id_clock_sync;
  int rc = 0;
  try {
#include "hordiff_itts_f2c.inc"      // generated: MOM_tracer_hor_diff.F90:368-387
  } catch (const std::string &) { rc = 1; }
  *num_itts_out = num_itts; *I_numitts_out = i_numitts;
  return rc;
}
