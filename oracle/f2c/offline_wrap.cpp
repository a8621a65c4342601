// TEST INFRASTRUCTURE ONLY -- the iteration loop of offline_advection_ale,
// /root/reference/src/tracer/MOM_offline_main.F90:270-415, and the body of remaining_transport_sum (:634-648), both
// UNMODIFIED and translated statement by statement by oracle/f2c/f2cpp.py --range.  This wrapper declares the
// routine's variables with the reference's declared shapes (:216-250), supplies the called routines of other modules
// -- advect_tracer is the translated reference itself (libref_f2c.so); pass_var / pass_vector are the oracle's
// group-pass emulation; reproducing_sum follows MOM_coms.F90:119-213,285 literally; the ALE routines abort (the tests
// keep MODULO(iter, off_ale_mod) /= 0); checksum / message / clock calls do nothing -- and includes the generated
// statements.  Same argument list as ora_offline_advection_loop (oracle/advect_ref.h).
#include "f2c_runtime.h"
#include "../advect_ref.h"
#include "obc_from_ora.h"
#include <cstdlib>

extern thread_local f2c_host *f2c_current_host;
// the translated reference routine (advect_f2c.inc, compiled into libref_f2c.so by ref_wrap.cpp)
void advect_tracer(Arr<double,3> h_end__, Arr<double,3> uhtr__, Arr<double,3> vhtr__, ocean_obc_type *obc, double dt, ocean_grid_type &g, verticalgrid_type &gv, unit_scale_type &us, tracer_advect_cs *cs, tracer_registry_type *reg, Opt<lbool> x_first_in, Arr<double,3> vol_prev__, Opt<int> max_iter_in, Opt<lbool> update_vol_prev, Arr<double,3> uhr_out__, Arr<double,3> vhr_out__);

namespace {
struct thermo_var_ptrs { Arr<double, 3> spv_avg; int valid_spv_halo = 0; };
struct offline_transport_cs {
  int num_off_iter = 0, off_ale_mod = 0;
  lbool debug = false, print_adv_offline = false;
  ocean_obc_type *obc = nullptr;
  double dt_offline = 0.0, min_residual = 0.0;
  tracer_advect_cs *tracer_adv_csp = nullptr;
  tracer_registry_type *tracer_reg = nullptr;
  void *ale_csp = nullptr;
  thermo_var_ptrs tv;
};

// keyword arguments of the advect_tracer call at :334-336, as tagged values
struct kw_x_first_in_t { lbool v; };        inline kw_x_first_in_t kw_x_first_in(lbool v) { return {v}; }
struct kw_vol_prev_t { Arr<double, 3> v; }; inline kw_vol_prev_t kw_vol_prev(Arr<double, 3> v) { return {v}; }
struct kw_max_iter_in_t { int v; };         inline kw_max_iter_in_t kw_max_iter_in(int v) { return {v}; }
struct kw_update_vol_prev_t { lbool v; };   inline kw_update_vol_prev_t kw_update_vol_prev(lbool v) { return {v}; }
struct kw_uhr_out_t { Arr<double, 3> v; };  inline kw_uhr_out_t kw_uhr_out(Arr<double, 3> v) { return {v}; }
struct kw_vhr_out_t { Arr<double, 3> v; };  inline kw_vhr_out_t kw_vhr_out(Arr<double, 3> v) { return {v}; }
inline double kw_scale(double v) { return v; }
inline double kw_dt(double v) { return v; }
void advect_tracer(Arr<double, 3> h_end, Arr<double, 3> uhtr, Arr<double, 3> vhtr, ocean_obc_type *obc, double dt,
                   ocean_grid_type &g, verticalgrid_type &gv, unit_scale_type &us, tracer_advect_cs *cs,
                   tracer_registry_type *reg, kw_x_first_in_t xf, kw_vol_prev_t vp, kw_max_iter_in_t mi,
                   kw_update_vol_prev_t uv, kw_uhr_out_t uo, kw_vhr_out_t vo) {
  ::advect_tracer(h_end, uhtr, vhtr, obc, dt, g, gv, us, cs, reg, Opt<lbool>(xf.v), vp.v, Opt<int>(mi.v), Opt<lbool>(uv.v), uo.v, vo.v);
}

// other modules' routines that only write messages / checksums / clocks
template <class... A> void hchksum(A &&...) {}
template <class... A> void uvchksum(A &&...) {}
template <class... A> void mom_tracer_chkinv(A &&...) {}
template <class... A> void mom_mesg(A &&...) {}
template <class... A> void cpu_clock_begin(A &&...) {}
template <class... A> void cpu_clock_end(A &&...) {}
// the ALE step of other modules (:346-380): not available here
[[noreturn]] void no_ale() { fprintf(stderr, "offline_wrap: the ALE step inside the loop is not translated\n"); abort(); }
template <class... A> void ale_update_regrid_weights(A &&...) { no_ale(); }
template <class... A> void pre_ale_adjustments(A &&...) { no_ale(); }
template <class... A> void ale_regrid(A &&...) { no_ale(); }
template <class... A> void ale_remap_tracers(A &&...) { no_ale(); }
inline int modulo(int a, int p) { int r = a % p; if (r != 0 && ((r < 0) != (p < 0))) r += p; return r; }

// pass_var / pass_vector on the oracle's single-rank emulation (FMS data movement, not part of the reference file)
struct PassCtx { ora_rank rank; const ora_domain *dom; std::vector<double> su, sv, sh; ora_tracer *Tr; };
PassCtx *g_pass = nullptr;
void pass_fields(double *u, double *v, double *h) {
  double *pu[1] = {u ? u : g_pass->su.data()}, *pv[1] = {v ? v : g_pass->sv.data()}, *ph[1] = {h ? h : g_pass->sh.data()};
  g_pass->rank.Tr = g_pass->Tr;
  ora_group_pass(1, &g_pass->rank, g_pass->dom, pu, pv, ph, 0);
}
template <class D> void pass_var(Arr<double, 3> h, D &&) { pass_fields(nullptr, nullptr, h.d); }
template <class D> void pass_vector(Arr<double, 3> u, Arr<double, 3> v, D &&) {
  pass_fields(u.d, v.d, nullptr);
  // symmetric memory: the boundary faces is-1 / js-1 take the neighbouring PE's (here: the periodic image's) value of
  // the same face; the group-pass emulation counts them as computational and leaves them, so it is done here
  const ora_grid &G = g_pass->rank.G;
  const ora_domain *dom = g_pass->dom;
  if (dom->symmetric) {
    if (dom->cyclic_x) for (int k = u.lo[2]; k <= u.hi[2]; k++) for (int j = G.jsd; j <= G.jed; j++) u(G.isc - 1, j, k) = u(G.iec, j, k);
    if (dom->cyclic_y) for (int k = v.lo[2]; k <= v.hi[2]; k++) for (int i = G.isd; i <= G.ied; i++) v(i, G.jsc - 1, k) = v(i, G.jec, k);
  }
}

// reproducing_sum_2d(array, isr, ier, jsr, jer), MOM_coms.F90:219-318 with reproducing_EFP_sum_2d :119-213
double reproducing_sum(Arr<double, 2> array, int isr, int ier, int jsr, int jer) {
  const long long prec = 1LL << 46;                                   // :27
  const double r_prec = 70368744177664.0, I_prec = 1.0 / 70368744177664.0;     // :28-29
  const int max_count_prec = (1 << (63 - 46)) - 1;                    // :30
  const double pr[6] = {r_prec * r_prec, r_prec, 1.0, 1.0 / r_prec, 1.0 / (r_prec * r_prec), 1.0 / (r_prec * r_prec * r_prec)};      // :36-37
  const double I_pr[6] = {1.0 / (r_prec * r_prec), 1.0 / r_prec, 1.0, r_prec, r_prec * r_prec, r_prec * r_prec * r_prec};            // :39-40
  const int lo0 = array.lo[0], lo1 = array.lo[1];      // "the array indices start at 1"
  long long ints_sum[6] = {0, 0, 0, 0, 0, 0};          // :154
  auto increment_ints_faster = [&](double r) {         // :597-626
    int sgn = 1; if (r < 0.0) sgn = -1;
    double rs = std::fabs(r);
    for (int i = 0; i < 6; i++) {
      long long ival = (long long)(rs * I_pr[i]);
      rs = rs - ival * pr[i];
      ints_sum[i] = ints_sum[i] + sgn * ival;
    }
  };
  auto carry_overflow = [&]() {                        // :629-649
    for (int i = 5; i >= 1; i--) if (std::llabs(ints_sum[i]) >= prec) {
      int num_carry = (int)(ints_sum[i] * I_prec);
      ints_sum[i] = ints_sum[i] - num_carry * prec;
      ints_sum[i - 1] = ints_sum[i - 1] + num_carry;
    }
  };
  const int is = isr, ie = ier, js = jsr, je = jer;
  if ((je + 1 - js) * (ie + 1 - is) < max_count_prec) {               // :156-160
    for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++) increment_ints_faster(array(i - 1 + lo0, j - 1 + lo1));
    carry_overflow();
  } else if ((ie + 1 - is) < max_count_prec) {                        // :161-167
    for (int j = js; j <= je; j++) {
      for (int i = is; i <= ie; i++) increment_ints_faster(array(i - 1 + lo0, j - 1 + lo1));
      carry_overflow();
    }
  } else { fprintf(stderr, "offline_wrap: rows longer than max_count_prec are not handled\n"); abort(); }
  // :195-211 sum_across_PEs (one PE) ; regularize_ints :653-691
  for (int i = 5; i >= 1; i--) if (std::llabs(ints_sum[i]) >= prec) {
    int num_carry = (int)(ints_sum[i] * I_prec);
    ints_sum[i] = ints_sum[i] - num_carry * prec;
    ints_sum[i - 1] = ints_sum[i - 1] + num_carry;
  }
  bool positive = true;
  for (int i = 0; i < 6; i++) if (std::llabs(ints_sum[i]) > 0) { if (ints_sum[i] < 0) positive = false; break; }
  if (positive) { for (int i = 5; i >= 1; i--) if (ints_sum[i] < 0) { ints_sum[i] = ints_sum[i] + prec; ints_sum[i - 1] = ints_sum[i - 1] - 1; } }
  else { for (int i = 5; i >= 1; i--) if (ints_sum[i] > 0) { ints_sum[i] = ints_sum[i] - prec; ints_sum[i - 1] = ints_sum[i - 1] + 1; } }
  double r = 0.0;                                                     // ints_to_real :550-559
  for (int i = 0; i < 6; i++) r = r + pr[i] * ints_sum[i];
  return r;
}

// real function remaining_transport_sum(G, GV, US, uhtr, vhtr, h_new), :615-650: declarations by hand, body generated
double remaining_transport_sum(ocean_grid_type &g, verticalgrid_type &gv, unit_scale_type &us, Arr<double, 3> uhtr,
                               Arr<double, 3> vhtr, Arr<double, 3> h_new) {
  Arr<double, 2> trans_rem_col = Arr<double, 2>::automatic({{g.isd, g.ied}, {g.jsd, g.jed}});      // :627
  double trans_cell = f_uninit<double>(), hl2_to_kg_scale = f_uninit<double>(), remaining_transport_sum = f_uninit<double>();
  int i = 0, j = 0, k = 0, is = 0, ie = 0, js = 0, je = 0, nz = 0;
#include "offline_rts_f2c.inc"      // generated: MOM_offline_main.F90:634-648
  return remaining_transport_sum;
}

// the group passes inside the translated advect_tracer (as in ref_wrap.cpp)
struct HostCtx { ora_tracer *Tr; int ntr; ora_rank rank; const ora_domain *dom; };
void host_pass(void *c, group_pass_type &g) {
  HostCtx *h = (HostCtx *)c;
  double *u[1] = {g.u}, *v[1] = {g.v}, *hp[1] = {g.scalars.at(0)};
  h->rank.Tr = h->Tr;
  ora_group_pass(1, &h->rank, h->dom, u, v, hp, h->ntr);
}
void host_sum(void *, int *, int) {}
}  // namespace

extern "C" int ref_offline_advection_loop(const ora_grid *G, const ora_cs *CS, ora_tracer *Tr, int ntr, double *h_pre_in,
                                          double *uhtr_in, double *vhtr_in, double dt_offline, int num_off_iter,
                                          int off_ale_mod, int x_before_y_in, double min_residual, double H_to_kg_m2,
                                          double L_to_m, double kg_m2_to_H, double m_to_L, const ora_domain *dom,
                                          const ora_obc *OBC, int *converged_out, int *n_iter_out, double *tot_residual_out,
                                          int nthreads) {
  (void)nthreads;
  if (!dom) return 1;
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
  gv.h_to_kg_m2 = H_to_kg_m2; gv.kg_m2_to_h = kg_m2_to_H;
  unit_scale_type us; us.l_to_m = L_to_m; us.m_to_l = m_to_L;
  tracer_advect_cs acs; acs.dt = CS->dt; acs.useppm = CS->usePPM != 0; acs.usehuynh = CS->useHuynh != 0;
  acs.usehuynhstencilbug = CS->useHuynhStencilBug != 0;
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
  offline_transport_cs cs_; offline_transport_cs *cs = &cs_;
  cs->num_off_iter = num_off_iter; cs->off_ale_mod = off_ale_mod; cs->dt_offline = dt_offline; cs->min_residual = min_residual;
  cs->tracer_adv_csp = &acs; cs->tracer_reg = &reg;
  ObcHolder obc_holder;
  cs->obc = obc_from_ora(OBC, obc_holder, G->nk);      // CS%OBC, handed to advect_tracer at :334
  // the routine's dummy arguments and locals with their declared shapes (:216-250)
  Arr<double, 3> h_pre = Arr<double, 3>::view(h_pre_in, {hi_, hj_, kk});
  Arr<double, 3> uhtr = Arr<double, 3>::view(uhtr_in, {ui_, hj_, kk}), vhtr = Arr<double, 3>::view(vhtr_in, {hi_, vj_, kk});
  Arr<double, 3> uhtr_sub = Arr<double, 3>::automatic({ui_, hj_, kk}), vhtr_sub = Arr<double, 3>::automatic({hi_, vj_, kk});
  Arr<double, 3> h_new = Arr<double, 3>::automatic({hi_, hj_, kk}), h_post_remap = Arr<double, 3>::automatic({hi_, hj_, kk});
  Arr<double, 3> h_vol = Arr<double, 3>::automatic({hi_, hj_, kk}), dzregrid = Arr<double, 3>::automatic({hi_, hj_, {1, G->nk + 1}});
  double prev_tot_residual = f_uninit<double>(), tot_residual = f_uninit<double>();
  int iter = 0, i = 0, j = 0, k = 0;
  const int is = G->isc, ie = G->iec, js = G->jsc, je = G->jec, nz = G->nk;                  // :257
  const int isd = G->isd, ied = G->ied, jsd = G->jsd, jed = G->jed;                          // :258
  const int isdb = G->IsdB, iedb = G->IedB, jsdb = G->JsdB, jedb = G->JedB;                  // :259
  lbool x_before_y = x_before_y_in != 0;                                                      // :254
  lbool converged = false;
  const double hl2_to_kg_scale = (L_to_m * L_to_m) * H_to_kg_m2;                              // :264 (messages only)
  const char *debug_msg = "", *mesg = ""; int id_clock_ale = 0;
  (void)hl2_to_kg_scale; (void)debug_msg; (void)mesg; (void)id_clock_ale; (void)i; (void)j; (void)k;
  std::vector<double> residuals;

  PassCtx pc; memset(&pc.rank, 0, sizeof pc.rank); pc.rank.G = *G; pc.rank.ioff = -(G->isc - 1); pc.rank.joff = -(G->jsc - 1);
  pc.dom = dom; pc.Tr = Tr;
  pc.su.assign((size_t)uhtr.size(), 0.0); pc.sv.assign((size_t)vhtr.size(), 0.0); pc.sh.assign((size_t)h_pre.size(), 0.0);
  g_pass = &pc;
  HostCtx hc; hc.Tr = Tr; hc.ntr = ntr; hc.dom = dom;
  memset(&hc.rank, 0, sizeof hc.rank); hc.rank.G = *G; hc.rank.ioff = -(G->isc - 1); hc.rank.joff = -(G->jsc - 1);
  f2c_host host; host.pass = host_pass; host.sum = host_sum; host.ctx = &hc;
  f2c_current_host = &host;
  int rc = 0;
  // remaining_transport_sum is recorded wherever the translated statements assign tot_residual
  struct Rec { std::vector<double> *v; double x; Rec &operator=(double y) { x = y; v->push_back(y); return *this; } operator double() const { return x; } };
  try {
#define tot_residual tot_residual_rec
    Rec tot_residual_rec{&residuals, tot_residual};
#include "offline_f2c.inc"          // generated: MOM_offline_main.F90:270-415
#undef tot_residual
  } catch (const std::string &) { rc = 1; }
  f2c_current_host = nullptr; g_pass = nullptr;
  if (converged_out) *converged_out = converged ? 1 : 0;
  if (n_iter_out) *n_iter_out = iter > num_off_iter ? num_off_iter : iter;
  if (tot_residual_out) for (size_t q = 0; q < residuals.size(); q++) tot_residual_out[q] = residuals[q];
  return rc;
}
