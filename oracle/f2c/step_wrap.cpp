// TEST INFRASTRUCTURE ONLY -- the loops either side of advect_tracer, UNMODIFIED reference
// lines translated statement by statement by oracle/f2c/f2cpp.py --range:
//   /root/reference/src/core/MOM_dynamics_split_RK2.F90:1045-1053     uhtr += uh*dt ; vhtr += vh*dt
//   /root/reference/src/tracer/MOM_tracer_registry.F90:744-767        frac_under_100m and khi
//   /root/reference/src/tracer/MOM_tracer_registry.F90:785-788, 795-798, 804-807   work2d of adv_xy_2d, zint, zint_100m
// The wrappers declare the routines' variables with the reference's shapes and include the generated statements.
#include "f2c_runtime.h"
#include "../advect_ref.h"

namespace {
template <class T, int N, class P>
bool f_any_section(const Arr<T, N> &a, const int (&fixed)[N], const bool (&full)[N], P pred) {
  static_assert(N == 3, "used for rank-3 arrays");
  for (int k = full[2] ? a.lo[2] : fixed[2]; k <= (full[2] ? a.hi[2] : fixed[2]); k++)
    for (int j = full[1] ? a.lo[1] : fixed[1]; j <= (full[1] ? a.hi[1] : fixed[1]); j++)
      for (int i = full[0] ? a.lo[0] : fixed[0]; i <= (full[0] ? a.hi[0] : fixed[0]); i++)
        if (pred(a(i, j, k))) return true;
  return false;
}
}  // namespace

// step_MOM_dyn_split_RK2: is..je are G%isc.., Isq = G%IscB, Ieq = G%IecB, Jsq = G%JscB, Jeq = G%JecB (:405-409)
extern "C" int ref_accumulate_transports(const ora_grid *G, double *uhtr_in, double *vhtr_in, const double *uh_in,
                                         const double *vh_in, double dt) {
  const Bnd hi_ = {G->isd, G->ied}, hj_ = {G->jsd, G->jed}, ui_ = {G->IsdB, G->IedB}, vj_ = {G->JsdB, G->JedB}, kk = {1, G->nk};
  Arr<double, 3> uhtr = Arr<double, 3>::view(uhtr_in, {ui_, hj_, kk}), vhtr = Arr<double, 3>::view(vhtr_in, {hi_, vj_, kk});
  Arr<double, 3> uh = Arr<double, 3>::view(const_cast<double *>(uh_in), {ui_, hj_, kk});
  Arr<double, 3> vh = Arr<double, 3>::view(const_cast<double *>(vh_in), {hi_, vj_, kk});
  const int is = G->isc, ie = G->iec, js = G->jsc, je = G->jec, nz = G->nk;
  const int isq = (G->IsdB == G->isd - 1) ? G->isc - 1 : G->isc, ieq = G->iec;      // MOM_hor_index.F90:87-96
  const int jsq = (G->JsdB == G->jsd - 1) ? G->jsc - 1 : G->jsc, jeq = G->jec;
  int i = 0, j = 0, k = 0;
  int rc = 0;
  try {
#include "step_acc_f2c.inc"      // generated: MOM_dynamics_split_RK2.F90:1045-1053
  } catch (const std::string &) { rc = 1; }
  return rc;
}

// post_tracer_transport_diagnostics: kind 0 adv_xy_2d (field3d = advection_xy), 1 zint, 2 zint_100m (field3d = Tr%t)
extern "C" int ref_column_diag(const ora_grid *G, int kind, const double *field3d, const double *h_diag_in, double H_to_m,
                               double *work2d_out, int *khi_out) {
  const Bnd hi_ = {G->isd, G->ied}, hj_ = {G->jsd, G->jed}, kk = {1, G->nk};
  verticalgrid_type gv; gv.ke = G->nk; gv.h_to_m = H_to_m;
  tracer_type tr;
  tr.t = Arr<double, 3>::view(const_cast<double *>(field3d), {hi_, hj_, kk});
  tr.advection_xy = Arr<double, 3>::view(const_cast<double *>(field3d), {hi_, hj_, kk});
  Arr<double, 3> h_diag = Arr<double, 3>::view(const_cast<double *>(h_diag_in), {hi_, hj_, kk});
  Arr<double, 2> work2d = Arr<double, 2>::view(work2d_out, {hi_, hj_});
  Arr<double, 3> frac_under_100m = Arr<double, 3>::automatic({hi_, hj_, kk});      // :735
  Arr<double, 2> ztop = Arr<double, 2>::automatic({hi_, hj_}), zbot = Arr<double, 2>::automatic({hi_, hj_});
  const int is = G->isc, ie = G->iec, js = G->jsc, je = G->jec, nz = G->nk;            // :741
  int i = 0, j = 0, k = 0, khi = 0;
  int rc = 0;
  try {
    if (kind == 0) {
#include "step_advxy2d_f2c.inc"  // generated: MOM_tracer_registry.F90:785-788
    } else if (kind == 1) {
#include "step_zint_f2c.inc"     // generated: :795-798
    } else {
      f_assign(frac_under_100m, 0.0);      // :744
      khi = 0;                             // :748
#include "step_frac_f2c.inc"     // generated: :752-767
#include "step_zint100_f2c.inc"  // generated: :804-807
    }
  } catch (const std::string &) { rc = 1; }
  if (khi_out) *khi_out = khi;
  return rc;
}
