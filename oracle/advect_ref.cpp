// TEST INFRASTRUCTURE ONLY -- CPU oracle for MOM6 advect_tracer.  Pinned bit for bit against a mechanical translation of the unmodified reference file (oracle/f2c, tests/test_f2c_reference.py); see advect_ref.h
// (see advect_ref.h).  Never linked, loaded or called by the product path.
//
// Statement-by-statement restatement of
//   /root/reference/src/tracer/MOM_tracer_advect.F90
//     advect_tracer :52-324, advect_x :329-701, advect_y :705-1087
// with gfortran semantics made explicit:
//   * MIN/MAX with 2-3 arguments are compare-select chains in argument order
//     (first argument wins ties, so +-0 ties resolve by argument order);
//   * SIGN(a,b) is copysign(a,b);
//   * parentheses honoured, same-precedence chains left to right;
//   * no FMA contraction: build with -ffp-contract=off (oracle/Makefile).
// The OBC branches (:442-478,580-628,825-862,965-1014) are restated too; OBC == NULL is the
// unassociated pointer.
//
// Uninitialised reads in the reference (SURVEY.md hazard 7): do_i outside the
// range written by the sweep is treated as .false. here.

#include "advect_ref.h"

#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
#include <limits>
#include <mutex>
#include <condition_variable>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

// ---- gfortran intrinsic semantics ------------------------------------------
inline double fmax2(double a, double b) { double m = a; if (b > m) m = b; return m; }
inline double fmin2(double a, double b) { double m = a; if (b < m) m = b; return m; }
inline double fmax3(double a, double b, double c) { double m = a; if (b > m) m = b; if (c > m) m = c; return m; }
inline double fmin3(double a, double b, double c) { double m = a; if (b < m) m = b; if (c < m) m = c; return m; }
inline double fsign(double a, double b) { return std::copysign(a, b); }

// ---- Fortran-indexed array views -------------------------------------------
struct V2 {
  double *p; int ilo, jlo; long ni;
  V2(const double *p_, int ilo_, int ihi_, int jlo_) : p(const_cast<double *>(p_)), ilo(ilo_), jlo(jlo_), ni(ihi_ - ilo_ + 1) {}
  inline double &operator()(int i, int j) const { return p[(long)(i - ilo) + ni * (long)(j - jlo)]; }
};
struct V3 {
  double *p; int ilo, jlo; long ni, nij;
  V3(const double *p_, int ilo_, int ihi_, int jlo_, int jhi_)
      : p(const_cast<double *>(p_)), ilo(ilo_), jlo(jlo_), ni(ihi_ - ilo_ + 1), nij((long)(ihi_ - ilo_ + 1) * (jhi_ - jlo_ + 1)) {}
  inline double &operator()(int i, int j, int k) const { return p[(long)(i - ilo) + ni * (long)(j - jlo) + nij * (long)(k - 1)]; }
  inline bool ok() const { return p != nullptr; }
};

struct Scratch {   // uninitialised work array, like a Fortran automatic array
  double *p;
  explicit Scratch(size_t n) : p((double *)std::malloc(sizeof(double) * (n ? n : 1))) {}
  ~Scratch() { std::free(p); }
  Scratch(const Scratch &) = delete;
  inline double &operator[](size_t q) const { return p[q]; }
};

struct Flags2 {  // logical (jlo:jhi, 1:nk)
  std::vector<unsigned char> v; int jlo; long nj;
  Flags2(int jlo_, int jhi_, int nk) : v((size_t)(jhi_ - jlo_ + 1) * nk, 0), jlo(jlo_), nj(jhi_ - jlo_ + 1) {}
  inline unsigned char &operator()(int j, int k) { return v[(size_t)(j - jlo) + nj * (k - 1)]; }
};

struct Ctx {
  const ora_grid &G;
  const ora_obc *OBC = nullptr;      // NULL == unassociated
  V2 areaT, IareaT, mask2dT, mask2dCu, mask2dCv;
  explicit Ctx(const ora_grid &g)
      : G(g), areaT(g.areaT, g.isd, g.ied, g.jsd), IareaT(g.IareaT, g.isd, g.ied, g.jsd),
        mask2dT(g.mask2dT, g.isd, g.ied, g.jsd), mask2dCu(g.mask2dCu, g.IsdB, g.IedB, g.jsd),
        mask2dCv(g.mask2dCv, g.isd, g.ied, g.JsdB) {}
};

// PLM slope, MOM_tracer_advect.F90:427-431 / :809-813.  m12 = mask*mask product
inline double plm_slope(double Tp, double Tc, double Tm, double mk1, double mk2) {
  double dMx = fmax3(Tp, Tc, Tm) - Tc;
  double dMn = Tc - fmin3(Tp, Tc, Tm);
  return mk1 * mk2 * fsign(fmin3(0.5 * std::fabs(Tp - Tm), 2.0 * dMx, 2.0 * dMn), Tp - Tm);
}

// ---------------------------------------------------------------------------
// advect_x, MOM_tracer_advect.F90:329-701 (one level k)
// ---------------------------------------------------------------------------
void advect_x(const Ctx &C, ora_tracer *Tr, int ntr, V3 &hprev, V3 &uhr, V2 &uh_neglect,
              Flags2 &domore_u, double Idt, int is, int ie, int js, int je, int k,
              bool usePPM, bool useHuynh) {
  const ora_grid &G = C.G;
  const int isd = G.isd, ied = G.ied, jsd = G.jsd, jed = G.jed, IsdB = G.IsdB, IedB = G.IedB;
  const long nI = IedB - IsdB + 1, nid = ied - isd + 1, njd = jed - jsd + 1;

  // work arrays (:355-371,386,390)
  Scratch slope_x_((size_t)nid * ntr), T_tmp_((size_t)nid * ntr);
  Scratch flux_x_((size_t)nI * njd * ntr);
  std::vector<double> uhh_(nI, 0.0), hlst_(nI, 0.0), Ihnew_(nI, 0.0), CFL_(nI, 0.0);
  std::vector<unsigned char> do_i_((size_t)nid * njd, 0);   // out-of-range reads == .false.
  std::vector<unsigned char> domore_u_initial(njd);
  auto slope_x = [&](int i, int m) -> double & { return slope_x_[(size_t)(i - isd) + nid * m]; };
  auto T_tmp = [&](int i, int m) -> double & { return T_tmp_[(size_t)(i - isd) + nid * m]; };
  auto flux_x = [&](int I, int j, int m) -> double & { return flux_x_[(size_t)(I - IsdB) + nI * ((size_t)(j - jsd) + njd * m)]; };
  auto uhh = [&](int I) -> double & { return uhh_[I - IsdB]; };
  auto hlst = [&](int I) -> double & { return hlst_[I - IsdB]; };
  auto Ihnew = [&](int I) -> double & { return Ihnew_[I - IsdB]; };
  auto CFL = [&](int I) -> double & { return CFL_[I - IsdB]; };
  auto do_i = [&](int i, int j) -> unsigned char & { return do_i_[(size_t)(i - isd) + nid * (j - jsd)]; };
  auto do_i_rd = [&](int i, int j) -> bool {
    if (i < isd || i > ied || j < jsd || j > jed) return false;
    return do_i_[(size_t)(i - isd) + nid * (j - jsd)] != 0;
  };

  // :394
  for (int j = jsd; j <= jed; j++) domore_u_initial[j - jsd] = domore_u(j, k);

  const bool usePLMslope = !(usePPM && useHuynh);   // :396
  int stencil = 1;                                   // :398
  if (usePPM && !useHuynh) stencil = 2;              // :399

  const double min_h = 0.1 * G.Angstrom_H;           // :401
  const double tiny_h = DBL_MIN;                     // :402 tiny(min_h)
  const double h_neglect = G.H_subroundoff;          // :403

  for (int I = is - 1; I <= ie; I++) CFL(I) = 0.0;   // :405

  for (int j = js; j <= je; j++) if (domore_u(j, k)) {   // :407
    domore_u(j, k) = 0;                                  // :408

    if (usePLMslope) {                                   // :411-433
      for (int m = 0; m < ntr; m++) {
        V3 t(Tr[m].t, isd, ied, jsd, jed);
        for (int i = is - stencil; i <= ie + stencil; i++) {
          double Tp = t(i + 1, j, k), Tc = t(i, j, k), Tm = t(i - 1, j, k);
          slope_x(i, m) = plm_slope(Tp, Tc, Tm, C.mask2dCu(i, j), C.mask2dCu(i - 1, j));
        }
      }
    }

    for (int m = 0; m < ntr; m++) {                      // :436-440
      V3 t(Tr[m].t, isd, ied, jsd, jed);
      for (int i = isd; i <= ied; i++) T_tmp(i, m) = t(i, j, k);
    }
    // loop through open boundaries and recalculate flux terms :442-478
    if (C.OBC) { if (C.OBC->OBC_pe) {
      for (int n = 0; n < C.OBC->number_of_segments; n++) {
        const ora_obc_segment *segment = &C.OBC->segment[n];
        if (segment->ntseg < 0) continue;                // .not. associated(segment%tr_Reg)
        if (segment->is_E_or_W) {
          if (j >= segment->jsd && j <= segment->jed) {
            const int I = segment->IsdB;                 // Fortran: I and i are one variable
            for (int m = 0; m < segment->ntseg; m++) {   // replace tracers with OBC values
              const int ntr_id = segment->Tr[m].ntr_index - 1;
              if (segment->Tr[m].tres) {
                V3 tres(segment->Tr[m].tres, segment->IsdB, segment->IedB, segment->jsd, segment->jed);
                if (segment->direction == ORA_OBC_DIRECTION_W) T_tmp(I, ntr_id) = tres(I, j, k);
                else T_tmp(I + 1, ntr_id) = tres(I, j, k);
              } else {
                if (segment->direction == ORA_OBC_DIRECTION_W) T_tmp(I, ntr_id) = segment->Tr[m].OBC_inflow_conc;
                else T_tmp(I + 1, ntr_id) = segment->Tr[m].OBC_inflow_conc;
              }
            }
            for (int m = 0; m < ntr; m++) {              // Apply update tracer values for slope calculation
              for (int i = segment->IsdB - 1; i <= segment->IsdB + 1; i++) {
                double Tp = T_tmp(i + 1, m), Tc = T_tmp(i, m), Tm = T_tmp(i - 1, m);
                slope_x(i, m) = plm_slope(Tp, Tc, Tm, C.mask2dCu(i, j), C.mask2dCu(i - 1, j));
              }
            }
          }
        }
      }
    } }

    for (int I = is - 1; I <= ie; I++) {                 // :485-514
      const int i = I;
      if ((uhr(I, j, k) == 0.0) ||
          ((uhr(I, j, k) < 0.0) && (hprev(i + 1, j, k) <= tiny_h)) ||
          ((uhr(I, j, k) > 0.0) && (hprev(i, j, k) <= tiny_h))) {
        uhh(I) = 0.0;
        CFL(I) = 0.0;
      } else if (uhr(I, j, k) < 0.0) {
        double hup = hprev(i + 1, j, k) - C.areaT(i + 1, j) * min_h;
        double hlos = fmax2(0.0, uhr(I + 1, j, k));
        if ((((hup - hlos) + uhr(I, j, k)) < 0.0) && ((0.5 * hup + uhr(I, j, k)) < 0.0)) {
          uhh(I) = fmin3(-0.5 * hup, -hup + hlos, 0.0);
          domore_u(j, k) = 1;
        } else {
          uhh(I) = uhr(I, j, k);
        }
        CFL(I) = -uhh(I) / (hprev(i + 1, j, k));
      } else {
        double hup = hprev(i, j, k) - C.areaT(i, j) * min_h;
        double hlos = fmax2(0.0, -uhr(I - 1, j, k));
        if ((((hup - hlos) - uhr(I, j, k)) < 0.0) && ((0.5 * hup - uhr(I, j, k)) < 0.0)) {
          uhh(I) = fmax3(0.5 * hup, hup - hlos, 0.0);
          domore_u(j, k) = 1;
        } else {
          uhh(I) = uhr(I, j, k);
        }
        CFL(I) = uhh(I) / (hprev(i, j, k));
      }
    }

    if (usePPM) {                                        // :517-557
      for (int m = 0; m < ntr; m++) for (int I = is - 1; I <= ie; I++) {
        const int i = I;
        int i_up;
        if (uhh(I) >= 0.0) i_up = i; else i_up = i + 1;
        double Tp = T_tmp(i_up + 1, m), Tc = T_tmp(i_up, m), Tm = T_tmp(i_up - 1, m);
        double aL, aR;
        if (useHuynh) {
          aL = (5. * Tc + (2. * Tm - Tp)) / 6.;
          aL = fmax2(fmin2(Tc, Tm), aL); aL = fmin2(fmax2(Tc, Tm), aL);
          aR = (5. * Tc + (2. * Tp - Tm)) / 6.;
          aR = fmax2(fmin2(Tc, Tp), aR); aR = fmin2(fmax2(Tc, Tp), aR);
        } else {
          aL = 0.5 * ((Tm + Tc) + (slope_x(i_up - 1, m) - slope_x(i_up, m)) / 3.);
          aR = 0.5 * ((Tc + Tp) + (slope_x(i_up, m) - slope_x(i_up + 1, m)) / 3.);
        }
        double dA = aR - aL, mA = 0.5 * (aR + aL);
        if (C.mask2dCu(i_up, j) * C.mask2dCu(i_up - 1, j) * (Tp - Tc) * (Tc - Tm) <= 0.) {
          aL = Tc; aR = Tc;
        } else if (dA * (Tc - mA) > (dA * dA) / 6.) {
          aL = 3. * Tc - 2. * aR;
        } else if (dA * (Tc - mA) < -(dA * dA) / 6.) {
          aR = 3. * Tc - 2. * aL;
        }
        double a6 = 6. * Tc - 3. * (aR + aL);
        if (uhh(I) >= 0.0) {
          flux_x(I, j, m) = uhh(I) * (aR - 0.5 * CFL(I) * ((aR - aL) - a6 * (1. - 2. / 3. * CFL(I))));
        } else {
          flux_x(I, j, m) = uhh(I) * (aL + 0.5 * CFL(I) * ((aR - aL) + a6 * (1. - 2. / 3. * CFL(I))));
        }
      }
    } else {                                             // PLM :559-577
      for (int m = 0; m < ntr; m++) for (int I = is - 1; I <= ie; I++) {
        const int i = I;
        if (uhh(I) >= 0.0) {
          double Tc = T_tmp(i, m);
          flux_x(I, j, m) = uhh(I) * (Tc + 0.5 * slope_x(i, m) * (1. - CFL(I)));
        } else {
          double Tc = T_tmp(i + 1, m);
          flux_x(I, j, m) = uhh(I) * (Tc - 0.5 * slope_x(i + 1, m) * (1. - CFL(I)));
        }
      }
    }
    if (C.OBC) { if (C.OBC->OBC_pe) {                     // :580-628
      if (C.OBC->specified_u_BCs_exist_globally || C.OBC->open_u_BCs_exist_globally) {
        for (int n = 0; n < C.OBC->number_of_segments; n++) {
          const ora_obc_segment *segment = &C.OBC->segment[n];
          if (segment->ntseg < 0) continue;
          if (segment->is_E_or_W) {
            if (j >= segment->jsd && j <= segment->jed) {
              const int I = segment->IsdB;
              // Tracer fluxes are set to prescribed values ... Now changing to simply fixed inflows.
              if (((uhr(I, j, k) > 0.0) && (segment->direction == ORA_OBC_DIRECTION_W)) ||
                  ((uhr(I, j, k) < 0.0) && (segment->direction == ORA_OBC_DIRECTION_E))) {
                uhh(I) = uhr(I, j, k);
                for (int m = 0; m < segment->ntseg; m++) {
                  const int ntr_id = segment->Tr[m].ntr_index - 1;
                  if (segment->Tr[m].tres) {
                    V3 tres(segment->Tr[m].tres, segment->IsdB, segment->IedB, segment->jsd, segment->jed);
                    flux_x(I, j, ntr_id) = uhh(I) * tres(I, j, k);
                  } else { flux_x(I, j, ntr_id) = uhh(I) * segment->Tr[m].OBC_inflow_conc; }
                }
              }
            }
          }
        }
      }
      if (C.OBC->open_u_BCs_exist_globally) {
        for (int n = 0; n < C.OBC->number_of_segments; n++) {
          const ora_obc_segment *segment = &C.OBC->segment[n];
          const int I = segment->IsdB, i = I;
          if (segment->is_E_or_W && (j >= segment->jsd && j <= segment->jed)) {
            if (segment->specified) continue;
            if (segment->ntseg < 0) continue;
            // Tracer fluxes are set to prescribed values only for inflows from masked areas.
            if (((uhr(I, j, k) > 0.0) && (C.mask2dT(i, j) < 0.5)) ||
                ((uhr(I, j, k) < 0.0) && (C.mask2dT(i + 1, j) < 0.5))) {
              uhh(I) = uhr(I, j, k);
              for (int m = 0; m < segment->ntseg; m++) {
                const int ntr_id = segment->Tr[m].ntr_index - 1;
                if (segment->Tr[m].tres) {
                  V3 tres(segment->Tr[m].tres, segment->IsdB, segment->IedB, segment->jsd, segment->jed);
                  flux_x(I, j, ntr_id) = uhh(I) * tres(I, j, k);
                } else { flux_x(I, j, ntr_id) = uhh(I) * segment->Tr[m].OBC_inflow_conc; }
              }
            }
          }
        }
      }
    } }

    for (int I = is - 1; I <= ie; I++) {                 // :632-635
      uhr(I, j, k) = uhr(I, j, k) - uhh(I);
      if (std::fabs(uhr(I, j, k)) < uh_neglect(I, j)) uhr(I, j, k) = 0.0;
    }
    for (int i = is; i <= ie; i++) {                     // :636-649
      const int I = i;
      if ((uhh(I) != 0.0) || (uhh(I - 1) != 0.0)) {
        do_i(i, j) = 1;
        hlst(i) = hprev(i, j, k);
        hprev(i, j, k) = hprev(i, j, k) - (uhh(I) - uhh(I - 1));
        if (hprev(i, j, k) <= 0.0) { do_i(i, j) = 0; }
        else if (hprev(i, j, k) < h_neglect * C.areaT(i, j)) {
          hlst(i) = hlst(i) + (h_neglect * C.areaT(i, j) - hprev(i, j, k));
          Ihnew(i) = 1.0 / (h_neglect * C.areaT(i, j));
        } else { Ihnew(i) = 1.0 / hprev(i, j, k); }
      } else {
        do_i(i, j) = 0;
      }
    }

    for (int m = 0; m < ntr; m++) {                      // :652-678
      V3 t(Tr[m].t, isd, ied, jsd, jed);
      for (int i = is; i <= ie; i++) {
        const int I = i;
        if (do_i(i, j)) {
          if (Ihnew(i) > 0.0) {
            t(i, j, k) = (t(i, j, k) * hlst(i) - (flux_x(I, j, m) - flux_x(I - 1, j, m))) * Ihnew(i);
          }
        }
      }
      if (Tr[m].ad_x) {                                  // :665-667
        V3 ad_x(Tr[m].ad_x, IsdB, IedB, jsd, jed);
        for (int I = is - 1; I <= ie; I++) { const int i = I;
          if (do_i_rd(i, j) || do_i_rd(i + 1, j)) ad_x(I, j, k) = ad_x(I, j, k) + flux_x(I, j, m) * Idt;
        }
      }
      if (Tr[m].advection_xy) {                          // :671-676
        V3 axy(Tr[m].advection_xy, isd, ied, jsd, jed);
        for (int i = is; i <= ie; i++) { const int I = i;
          if (do_i(i, j)) axy(i, j, k) = axy(i, j, k) - (flux_x(I, j, m) - flux_x(I - 1, j, m)) * Idt * C.IareaT(i, j);
        }
      }
    }
  }  // j loop :680

  for (int m = 0; m < ntr; m++) if (Tr[m].conc_underflow > 0.0) {   // :683-687
    V3 t(Tr[m].t, isd, ied, jsd, jed);
    for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++)
      if (std::fabs(t(i, j, k)) < Tr[m].conc_underflow) t(i, j, k) = 0.0;
  }

  // :691-699 (ordered over k: serialised so the k-sum keeps the reference order)
#pragma omp ordered
  {
    for (int m = 0; m < ntr; m++) if (Tr[m].ad2d_x) {
      V2 ad2d_x(Tr[m].ad2d_x, IsdB, IedB, jsd);
      for (int j = js; j <= je; j++) if (domore_u_initial[j - jsd]) {
        for (int I = is - 1; I <= ie; I++) { const int i = I;
          if (do_i_rd(i, j) || do_i_rd(i + 1, j)) ad2d_x(I, j) = ad2d_x(I, j) + flux_x(I, j, m) * Idt;
        }
      }
    }
  }
}

// ---------------------------------------------------------------------------
// advect_y, MOM_tracer_advect.F90:705-1087 (one level k)
// ---------------------------------------------------------------------------
void advect_y(const Ctx &C, ora_tracer *Tr, int ntr, V3 &hprev, V3 &vhr, V2 &vh_neglect,
              Flags2 &domore_v, double Idt, int is, int ie, int js, int je, int k,
              bool usePPM, bool useHuynh) {
  const ora_grid &G = C.G;
  const int isd = G.isd, ied = G.ied, jsd = G.jsd, jed = G.jed, JsdB = G.JsdB, JedB = G.JedB;
  const long nid = ied - isd + 1, njd = jed - jsd + 1, nJ = JedB - JsdB + 1;

  // work arrays (:731-766).  slope_y(i,m,j), flux_y(i,m,J), T_tmp(i,m,j)
  Scratch slope_y_((size_t)nid * ntr * njd);
  Scratch flux_y_((size_t)nid * ntr * nJ);
  Scratch T_tmp_((size_t)nid * ntr * nJ);
  std::vector<double> vhh_((size_t)nid * nJ, 0.0);
  std::vector<double> hlst_(nid + 1, 0.0), Ihnew_(nid + 1, 0.0), CFL_(nid + 1, 0.0);
  // do_j_tr(SZJ_): the reference indexes j+j2 with j in js-1..je, j2 in 1-stencil..stencil
  std::vector<unsigned char> do_j_tr_(njd + 8, 0);
  std::vector<unsigned char> do_i_((size_t)nid * njd, 0);
  std::vector<unsigned char> domore_v_initial(nJ);
  auto slope_y = [&](int i, int m, int j) -> double & { return slope_y_[(size_t)(i - isd) + nid * ((size_t)m + ntr * (size_t)(j - jsd))]; };
  auto flux_y = [&](int i, int m, int J) -> double & { return flux_y_[(size_t)(i - isd) + nid * ((size_t)m + ntr * (size_t)(J - JsdB))]; };
  auto T_tmp = [&](int i, int m, int j) -> double & { return T_tmp_[(size_t)(i - isd) + nid * ((size_t)m + ntr * (size_t)(j - JsdB))]; };
  auto vhh = [&](int i, int J) -> double & { return vhh_[(size_t)(i - isd) + nid * (size_t)(J - JsdB)]; };
  auto hlst = [&](int i) -> double & { return hlst_[i - isd]; };
  auto Ihnew = [&](int i) -> double & { return Ihnew_[i - isd]; };
  auto CFL = [&](int i) -> double & { return CFL_[i - isd]; };
  auto do_j_tr = [&](int j) -> unsigned char & { return do_j_tr_[j - jsd + 4]; };
  auto do_i = [&](int i, int j) -> unsigned char & { return do_i_[(size_t)(i - isd) + nid * (j - jsd)]; };
  auto do_i_rd = [&](int i, int j) -> bool {
    if (i < isd || i > ied || j < jsd || j > jed) return false;
    return do_i_[(size_t)(i - isd) + nid * (j - jsd)] != 0;
  };

  const bool usePLMslope = !(usePPM && useHuynh);   // :768
  int stencil = 1;                                   // :770
  if (usePPM && !useHuynh) stencil = 2;              // :771

  const double min_h = 0.1 * G.Angstrom_H;           // :773
  const double tiny_h = DBL_MIN;                     // :774
  const double h_neglect = G.H_subroundoff;          // :775

  // :787-789
  for (int J = js - 1; J <= je; J++) if (domore_v(J, k)) {
    for (int j2 = 1 - stencil; j2 <= stencil; j2++) do_j_tr(J + j2) = 1;
  }
  for (int J = JsdB; J <= JedB; J++) domore_v_initial[J - JsdB] = domore_v(J, k);

  if (usePLMslope) {                                 // :793-815
    for (int j = js - stencil; j <= je + stencil; j++) if (do_j_tr(j)) {
      for (int m = 0; m < ntr; m++) {
        V3 t(Tr[m].t, isd, ied, jsd, jed);
        for (int i = is; i <= ie; i++) {
          double Tp = t(i, j + 1, k), Tc = t(i, j, k), Tm = t(i, j - 1, k);
          slope_y(i, m, j) = plm_slope(Tp, Tc, Tm, C.mask2dCv(i, j), C.mask2dCv(i, j - 1));
        }
      }
    }
  }

  for (int j = jsd; j <= jed; j++) for (int m = 0; m < ntr; m++) {   // :820-822
    V3 t(Tr[m].t, isd, ied, jsd, jed);
    for (int i = isd; i <= ied; i++) T_tmp(i, m, j) = t(i, j, k);
  }
  // loop through open boundaries and recalculate flux terms :825-862
  if (C.OBC) { if (C.OBC->OBC_pe) {
    for (int n = 0; n < C.OBC->number_of_segments; n++) {
      const ora_obc_segment *segment = &C.OBC->segment[n];
      if (segment->ntseg < 0) continue;
      for (int i = is; i <= ie; i++) {
        if (segment->is_N_or_S) {
          if (i >= segment->isd && i <= segment->ied) {
            const int J = segment->JsdB;                 // Fortran: J and j are one variable
            for (int m = 0; m < segment->ntseg; m++) {   // replace tracers with OBC values
              const int ntr_id = segment->Tr[m].ntr_index - 1;
              if (segment->Tr[m].tres) {
                V3 tres(segment->Tr[m].tres, segment->isd, segment->ied, segment->JsdB, segment->JedB);
                if (segment->direction == ORA_OBC_DIRECTION_S) T_tmp(i, ntr_id, J) = tres(i, J, k);
                else T_tmp(i, ntr_id, J + 1) = tres(i, J, k);
              } else {
                if (segment->direction == ORA_OBC_DIRECTION_S) T_tmp(i, ntr_id, J) = segment->Tr[m].OBC_inflow_conc;
                else T_tmp(i, ntr_id, J + 1) = segment->Tr[m].OBC_inflow_conc;
              }
            }
            for (int m = 0; m < ntr; m++) {              // Apply update tracer values for slope calculation
              for (int j = segment->JsdB - 1; j <= segment->JsdB + 1; j++) {
                double Tp = T_tmp(i, m, j + 1), Tc = T_tmp(i, m, j), Tm = T_tmp(i, m, j - 1);
                slope_y(i, m, j) = plm_slope(Tp, Tc, Tm, C.mask2dCv(i, j), C.mask2dCv(i, j - 1));
              }
            }
          }
        }
      }
    }
  } }

  for (int J = js - 1; J <= je; J++) {
    const int j = J;
    if (domore_v(J, k)) {                             // :868
      domore_v(J, k) = 0;                             // :869
      for (int i = is; i <= ie; i++) {                // :871-900
        if ((vhr(i, J, k) == 0.0) ||
            ((vhr(i, J, k) < 0.0) && (hprev(i, j + 1, k) <= tiny_h)) ||
            ((vhr(i, J, k) > 0.0) && (hprev(i, j, k) <= tiny_h))) {
          vhh(i, J) = 0.0;
          CFL(i) = 0.0;
        } else if (vhr(i, J, k) < 0.0) {
          double hup = hprev(i, j + 1, k) - C.areaT(i, j + 1) * min_h;
          double hlos = fmax2(0.0, vhr(i, J + 1, k));
          if ((((hup - hlos) + vhr(i, J, k)) < 0.0) && ((0.5 * hup + vhr(i, J, k)) < 0.0)) {
            vhh(i, J) = fmin3(-0.5 * hup, -hup + hlos, 0.0);
            domore_v(J, k) = 1;
          } else {
            vhh(i, J) = vhr(i, J, k);
          }
          CFL(i) = -vhh(i, J) / hprev(i, j + 1, k);
        } else {
          double hup = hprev(i, j, k) - C.areaT(i, j) * min_h;
          double hlos = fmax2(0.0, -vhr(i, J - 1, k));
          if ((((hup - hlos) - vhr(i, J, k)) < 0.0) && ((0.5 * hup - vhr(i, J, k)) < 0.0)) {
            vhh(i, J) = fmax3(0.5 * hup, hup - hlos, 0.0);
            domore_v(J, k) = 1;
          } else {
            vhh(i, J) = vhr(i, J, k);
          }
          CFL(i) = vhh(i, J) / hprev(i, j, k);
        }
      }

      if (usePPM) {                                   // :902-942
        for (int m = 0; m < ntr; m++) for (int i = is; i <= ie; i++) {
          int j_up;
          if (vhh(i, J) >= 0.0) j_up = j; else j_up = j + 1;
          double Tp = T_tmp(i, m, j_up + 1), Tc = T_tmp(i, m, j_up), Tm = T_tmp(i, m, j_up - 1);
          double aL, aR;
          if (useHuynh) {
            aL = (5. * Tc + (2. * Tm - Tp)) / 6.;
            aL = fmax2(fmin2(Tc, Tm), aL); aL = fmin2(fmax2(Tc, Tm), aL);
            aR = (5. * Tc + (2. * Tp - Tm)) / 6.;
            aR = fmax2(fmin2(Tc, Tp), aR); aR = fmin2(fmax2(Tc, Tp), aR);
          } else {
            aL = 0.5 * ((Tm + Tc) + (slope_y(i, m, j_up - 1) - slope_y(i, m, j_up)) / 3.);
            aR = 0.5 * ((Tc + Tp) + (slope_y(i, m, j_up) - slope_y(i, m, j_up + 1)) / 3.);
          }
          double dA = aR - aL, mA = 0.5 * (aR + aL);
          if (C.mask2dCv(i, j_up) * C.mask2dCv(i, j_up - 1) * (Tp - Tc) * (Tc - Tm) <= 0.) {
            aL = Tc; aR = Tc;
          } else if (dA * (Tc - mA) > (dA * dA) / 6.) {
            aL = 3. * Tc - 2. * aR;
          } else if (dA * (Tc - mA) < -(dA * dA) / 6.) {
            aR = 3. * Tc - 2. * aL;
          }
          double a6 = 6. * Tc - 3. * (aR + aL);
          if (vhh(i, J) >= 0.0) {
            flux_y(i, m, J) = vhh(i, J) * (aR - 0.5 * CFL(i) * ((aR - aL) - a6 * (1. - 2. / 3. * CFL(i))));
          } else {
            flux_y(i, m, J) = vhh(i, J) * (aL + 0.5 * CFL(i) * ((aR - aL) + a6 * (1. - 2. / 3. * CFL(i))));
          }
        }
      } else {                                        // PLM :943-963
        for (int m = 0; m < ntr; m++) for (int i = is; i <= ie; i++) {
          if (vhh(i, J) >= 0.0) {
            double Tc = T_tmp(i, m, j);
            flux_y(i, m, J) = vhh(i, J) * (Tc + 0.5 * slope_y(i, m, j) * (1. - CFL(i)));
          } else {
            double Tc = T_tmp(i, m, j + 1);
            flux_y(i, m, J) = vhh(i, J) * (Tc - 0.5 * slope_y(i, m, j + 1) * (1. - CFL(i)));
          }
        }
      }
      if (C.OBC) { if (C.OBC->OBC_pe) {                 // :965-1014
        if (C.OBC->specified_v_BCs_exist_globally || C.OBC->open_v_BCs_exist_globally) {
          for (int n = 0; n < C.OBC->number_of_segments; n++) {
            const ora_obc_segment *segment = &C.OBC->segment[n];
            if (!segment->specified) continue;
            if (segment->ntseg < 0) continue;
            if (segment->is_N_or_S) {
              if (J >= segment->JsdB && J <= segment->JedB) {
                for (int i = segment->isd; i <= segment->ied; i++) {
                  if (((vhr(i, J, k) > 0.0) && (segment->direction == ORA_OBC_DIRECTION_S)) ||
                      ((vhr(i, J, k) < 0.0) && (segment->direction == ORA_OBC_DIRECTION_N))) {
                    vhh(i, J) = vhr(i, J, k);
                    for (int m = 0; m < segment->ntseg; m++) {
                      const int ntr_id = segment->Tr[m].ntr_index - 1;
                      if (segment->Tr[m].tres) {
                        V3 tres(segment->Tr[m].tres, segment->isd, segment->ied, segment->JsdB, segment->JedB);
                        flux_y(i, ntr_id, J) = vhh(i, J) * tres(i, J, k);
                      } else { flux_y(i, ntr_id, J) = vhh(i, J) * segment->Tr[m].OBC_inflow_conc; }
                    }
                  }
                }
              }
            }
          }
        }
        if (C.OBC->open_v_BCs_exist_globally) {
          for (int n = 0; n < C.OBC->number_of_segments; n++) {
            const ora_obc_segment *segment = &C.OBC->segment[n];
            if (segment->specified) continue;
            if (segment->ntseg < 0) continue;
            if (segment->is_N_or_S && (J >= segment->JsdB && J <= segment->JedB)) {
              for (int i = segment->isd; i <= segment->ied; i++) {
                // Tracer fluxes are set to prescribed values only for inflows from masked areas.
                if (((vhr(i, J, k) > 0.0) && (C.mask2dT(i, j) < 0.5)) ||
                    ((vhr(i, J, k) < 0.0) && (C.mask2dT(i, j + 1) < 0.5))) {
                  vhh(i, J) = vhr(i, J, k);
                  for (int m = 0; m < segment->ntseg; m++) {
                    const int ntr_id = segment->Tr[m].ntr_index - 1;
                    if (segment->Tr[m].tres) {
                      V3 tres(segment->Tr[m].tres, segment->isd, segment->ied, segment->JsdB, segment->JedB);
                      flux_y(i, ntr_id, J) = vhh(i, J) * tres(i, J, k);
                    } else { flux_y(i, ntr_id, J) = vhh(i, J) * segment->Tr[m].OBC_inflow_conc; }
                  }
                }
              }
            }
          }
        }
      } }
    } else {                                          // :1016-1018
      for (int i = is; i <= ie; i++) vhh(i, J) = 0.0;
      for (int m = 0; m < ntr; m++) for (int i = is; i <= ie; i++) flux_y(i, m, J) = 0.0;
    }
  }

  for (int J = js - 1; J <= je; J++) for (int i = is; i <= ie; i++) {   // :1021-1024
    vhr(i, J, k) = vhr(i, J, k) - vhh(i, J);
    if (std::fabs(vhr(i, J, k)) < vh_neglect(i, J)) vhr(i, J, k) = 0.0;
  }

  for (int j = js; j <= je; j++) if (do_j_tr(j)) {    // :1028-1059
    const int J = j;
    for (int i = is; i <= ie; i++) {
      if ((vhh(i, J) != 0.0) || (vhh(i, J - 1) != 0.0)) {
        do_i(i, j) = 1;
        hlst(i) = hprev(i, j, k);
        hprev(i, j, k) = fmax2(hprev(i, j, k) - (vhh(i, J) - vhh(i, J - 1)), 0.0);
        if (hprev(i, j, k) <= 0.0) { do_i(i, j) = 0; }
        else if (hprev(i, j, k) < h_neglect * C.areaT(i, j)) {
          hlst(i) = hlst(i) + (h_neglect * C.areaT(i, j) - hprev(i, j, k));
          Ihnew(i) = 1.0 / (h_neglect * C.areaT(i, j));
        } else { Ihnew(i) = 1.0 / hprev(i, j, k); }
      } else { do_i(i, j) = 0; }
    }
    for (int m = 0; m < ntr; m++) {
      V3 t(Tr[m].t, isd, ied, jsd, jed);
      for (int i = is; i <= ie; i++) if (do_i(i, j)) {
        t(i, j, k) = (t(i, j, k) * hlst(i) - (flux_y(i, m, J) - flux_y(i, m, J - 1))) * Ihnew(i);
      }
      if (Tr[m].advection_xy) {
        V3 axy(Tr[m].advection_xy, isd, ied, jsd, jed);
        for (int i = is; i <= ie; i++) if (do_i(i, j)) {
          axy(i, j, k) = axy(i, j, k) - (flux_y(i, m, J) - flux_y(i, m, J - 1)) * Idt * C.IareaT(i, j);
        }
      }
    }
  }

  for (int m = 0; m < ntr; m++) if (Tr[m].conc_underflow > 0.0) {   // :1062-1066
    V3 t(Tr[m].t, isd, ied, jsd, jed);
    for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++)
      if (std::fabs(t(i, j, k)) < Tr[m].conc_underflow) t(i, j, k) = 0.0;
  }

  // :1069-1085
#pragma omp ordered
  {
    for (int m = 0; m < ntr; m++) if (Tr[m].ad_y) {
      V3 ad_y(Tr[m].ad_y, isd, ied, JsdB, JedB);
      for (int J = js - 1; J <= je; J++) if (domore_v_initial[J - JsdB]) { const int j = J;
        for (int i = is; i <= ie; i++) if (do_i_rd(i, j) || do_i_rd(i, j + 1))
          ad_y(i, J, k) = ad_y(i, J, k) + flux_y(i, m, J) * Idt;
      }
    }
    for (int m = 0; m < ntr; m++) if (Tr[m].ad2d_y) {
      V2 ad2d_y(Tr[m].ad2d_y, isd, ied, JsdB);
      for (int J = js - 1; J <= je; J++) if (domore_v_initial[J - JsdB]) { const int j = J;
        for (int i = is; i <= ie; i++) if (do_i_rd(i, j) || do_i_rd(i, j + 1))
          ad2d_y(i, J) = ad2d_y(i, J) + flux_y(i, m, J) * Idt;
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Emulated halo exchange (stand-in for FMS mpp_do_group_update as used at
// MOM_tracer_advect.F90:140-144,206 through
// config_src/infra/FMS2/MOM_domain_infra.F90:983-1025,1084-1137,1141-1158):
// full-width halos, all sides and corners, CGRID_NE vector pair (uhr,vhr) with
// sign change across a tripolar north fold, scalars hprev and every tracer.
// Points with no source (closed boundaries) are left untouched, as FMS does.
// ---------------------------------------------------------------------------
enum Pos { POS_H = 0, POS_U = 1, POS_V = 2 };

struct Src { int rank; int i, j; double sgn; };

// Locate the owner of global point (ig,jg) of position pos.  Returns false if
// the point lies outside the global domain (closed boundary).
bool locate(int nranks, const ora_rank *ranks, const ora_domain *dom, Pos pos, int ig, int jg, Src *s) {
  const int nig = dom->nig, njg = dom->njg;
  double sgn = 1.0;
  // h point (ig,jg) is cell ig in 1..nig.  u point Ig is the east face of cell
  // Ig (Ig in 0..nig); v point Jg is the north face of cell Jg (Jg in 0..njg).
  // Tripolar north fold: cells above row njg map to (nig+1-ig, 2*njg+1-jg).
  // The fold line is the v-face row Jg == njg (north edge of the last cell row).
  if (dom->tripolar_n && jg > njg) {
    if (pos == POS_H) { ig = nig + 1 - ig; jg = 2 * njg + 1 - jg; }
    else if (pos == POS_U) { ig = nig - ig; jg = 2 * njg + 1 - jg; sgn = -1.0; }
    else { ig = nig + 1 - ig; jg = 2 * njg - jg; sgn = -1.0; }
  }
  const int ilo = (pos == POS_U) ? 0 : 1, jlo = (pos == POS_V) ? 0 : 1;
  if (dom->cyclic_x) {
    // u faces 0 and nig coincide; map into 1..nig (or keep 0 only if it exists nowhere else)
    while (ig < 1) ig += nig;
    while (ig > nig) ig -= nig;
  } else if (ig < ilo || ig > nig) return false;
  if (dom->cyclic_y) {
    while (jg < 1) jg += njg;
    while (jg > njg) jg -= njg;
  } else if (jg < jlo || jg > njg) return false;
  for (int r = 0; r < nranks; r++) {
    const ora_grid &G = ranks[r].G;
    int i = ig - ranks[r].ioff, j = jg - ranks[r].joff;
    int i0 = G.isc, j0 = G.jsc;
    if (pos == POS_U && ig == 0) i0 = G.isc - 1;   // western domain-boundary face
    if (pos == POS_V && jg == 0) j0 = G.jsc - 1;   // southern domain-boundary face
    if (i >= i0 && i <= G.iec && j >= j0 && j <= G.jec) { s->rank = r; s->i = i; s->j = j; s->sgn = sgn; return true; }
  }
  return false;
}

// Is local point (i,j) of position pos inside rank G's own (non-halo) range?
inline bool in_compute(const ora_grid &G, const ora_domain *dom, Pos pos, int i, int j) {
  int i0 = G.isc, j0 = G.jsc;
  if (dom->symmetric) { if (pos == POS_U) i0 = G.isc - 1; if (pos == POS_V) j0 = G.jsc - 1; }
  return i >= i0 && i <= G.iec && j >= j0 && j <= G.jec;
}

struct HaloItem { long dst; int srank; long src; double sgn; };

void build_halo_list(int nranks, const ora_rank *ranks, const ora_domain *dom, int r, Pos pos, std::vector<HaloItem> &out) {
  const ora_grid &G = ranks[r].G;
  const int ilo = (pos == POS_U) ? G.IsdB : G.isd, jlo = (pos == POS_V) ? G.JsdB : G.jsd;
  const long ni = G.ied - ilo + 1;
  for (int j = jlo; j <= G.jed; j++) for (int i = ilo; i <= G.ied; i++) {
    if (in_compute(G, dom, pos, i, j)) continue;
    Src s;
    if (!locate(nranks, ranks, dom, pos, i + ranks[r].ioff, j + ranks[r].joff, &s)) continue;
    const ora_grid &Gs = ranks[s.rank].G;
    const int silo = (pos == POS_U) ? Gs.IsdB : Gs.isd, sjlo = (pos == POS_V) ? Gs.JsdB : Gs.jsd;
    const long sni = Gs.ied - silo + 1;
    HaloItem it;
    it.dst = (long)(i - ilo) + ni * (long)(j - jlo);
    it.srank = s.rank;
    it.src = (long)(s.i - silo) + sni * (long)(s.j - sjlo);
    it.sgn = s.sgn;
    out.push_back(it);
  }
}

long plane_size(const ora_grid &G, Pos pos) {
  const long ni = G.ied - ((pos == POS_U) ? G.IsdB : G.isd) + 1;
  const long nj = G.jed - ((pos == POS_V) ? G.JsdB : G.jsd) + 1;
  return ni * nj;
}

// Apply the exchange to rank r only (sources are other ranks' compute points,
// which nobody writes during the exchange, so ranks may run this concurrently).
void group_pass_rank(int nranks, ora_rank *ranks, const ora_domain *dom, int r,
                     double **uhr, double **vhr, double **hprev, int ntr) {
  const int nk = ranks[r].G.nk;
  for (int p = 0; p < 3; p++) {
    Pos pos = (Pos)p;
    std::vector<HaloItem> list;
    build_halo_list(nranks, ranks, dom, r, pos, list);
    const long psz = plane_size(ranks[r].G, pos);
    auto apply = [&](double *dst, double **srcs_of_rank, bool vec) {
      for (int k = 0; k < nk; k++) for (const HaloItem &it : list) {
        const long spsz = plane_size(ranks[it.srank].G, pos);
        double v = srcs_of_rank[it.srank][it.src + spsz * k];
        dst[it.dst + psz * k] = (vec && it.sgn < 0.0) ? -v : v;
      }
    };
    if (pos == POS_U) apply(uhr[r], uhr, true);
    else if (pos == POS_V) apply(vhr[r], vhr, true);
    else {
      apply(hprev[r], hprev, false);
      std::vector<double *> tp(nranks);
      for (int m = 0; m < ntr; m++) {
        for (int q = 0; q < nranks; q++) tp[q] = ranks[q].Tr[m].t;
        apply(ranks[r].Tr[m].t, tp.data(), false);
      }
    }
  }
}

struct SingleCtx {   // built-in single-rank pass
  ora_rank rank; const ora_domain *dom; int ntr;
};
void single_pass(void *ctx, double *uhr, double *vhr, double *hprev, ora_tracer *Tr, int ntr) {
  SingleCtx *s = (SingleCtx *)ctx;
  s->rank.Tr = Tr;
  double *u[1] = {uhr}, *v[1] = {vhr}, *h[1] = {hprev};
  // A halo point whose source is this rank's own compute point: in-place is
  // safe because sources (compute points) are never written by the exchange.
  group_pass_rank(1, &s->rank, s->dom, 0, u, v, h, ntr);
}
void single_sum(void *, int *, int) {}

// Lock-step barrier for the multi-rank emulation.
struct Barrier {
  std::mutex m; std::condition_variable cv; int n, count = 0, gen = 0;
  explicit Barrier(int n_) : n(n_) {}
  void wait() {
    std::unique_lock<std::mutex> lk(m);
    int g = gen;
    if (++count == n) { count = 0; gen++; cv.notify_all(); }
    else cv.wait(lk, [&] { return g != gen; });
  }
};
struct MultiShared {
  int nranks; ora_rank *ranks; const ora_domain *dom; int ntr;
  Barrier bar;
  std::vector<double *> uhr, vhr, hprev;
  std::vector<int> ksum;
  MultiShared(int n) : bar(n), uhr(n), vhr(n), hprev(n) {}
};
struct MultiCtx { MultiShared *sh; int r; };
void multi_pass(void *ctx, double *uhr, double *vhr, double *hprev, ora_tracer *, int ntr) {
  MultiCtx *c = (MultiCtx *)ctx; MultiShared *sh = c->sh;
  sh->uhr[c->r] = uhr; sh->vhr[c->r] = vhr; sh->hprev[c->r] = hprev;
  sh->bar.wait();
  group_pass_rank(sh->nranks, sh->ranks, sh->dom, c->r, sh->uhr.data(), sh->vhr.data(), sh->hprev.data(), ntr);
  sh->bar.wait();
}
void multi_sum(void *ctx, int *domore_k, int nz) {
  MultiCtx *c = (MultiCtx *)ctx; MultiShared *sh = c->sh;
  if (c->r == 0) sh->ksum.assign(nz, 0);
  sh->bar.wait();
  { std::lock_guard<std::mutex> lk(sh->bar.m); for (int k = 0; k < nz; k++) sh->ksum[k] += domore_k[k]; }
  sh->bar.wait();
  for (int k = 0; k < nz; k++) domore_k[k] = sh->ksum[k];
  sh->bar.wait();
}

}  // namespace

// ---------------------------------------------------------------------------
// advect_tracer, MOM_tracer_advect.F90:52-324
// ---------------------------------------------------------------------------
extern "C" int ora_advect_tracer(const ora_grid *Gp, const ora_cs *CS, ora_tracer *Tr, int ntr,
                                 const double *h_end_, const double *uhtr_, const double *vhtr_,
                                 double dt, int x_first_in, double *vol_prev_, int max_iter_in,
                                 int update_vol_prev, double *uhr_out, double *vhr_out,
                                 const ora_domain *dom, ora_pass_fn pass, ora_sum_fn sum,
                                 void *ctx, int *n_iter_out, int nthreads) {
  return ora_advect_tracer_obc(Gp, CS, Tr, ntr, h_end_, uhtr_, vhtr_, dt, x_first_in, vol_prev_, max_iter_in,
                               update_vol_prev, uhr_out, vhr_out, dom, pass, sum, ctx, n_iter_out, nthreads, nullptr);
}

extern "C" int ora_advect_tracer_obc(const ora_grid *Gp, const ora_cs *CS, ora_tracer *Tr, int ntr,
                                     const double *h_end_, const double *uhtr_, const double *vhtr_,
                                     double dt, int x_first_in, double *vol_prev_, int max_iter_in,
                                     int update_vol_prev, double *uhr_out, double *vhr_out,
                                     const ora_domain *dom, ora_pass_fn pass, ora_sum_fn sum,
                                     void *ctx, int *n_iter_out, int nthreads, const ora_obc *OBC) {
  if (n_iter_out) *n_iter_out = 0;
  if (!CS || !Gp) return 1;               // :120-121
  if (!Tr && ntr != 0) return 1;          // :122-123
  if (ntr == 0) return 0;                 // :124
  const ora_grid &G = *Gp;
  Ctx C(G);
  C.OBC = OBC;
  (void)nthreads;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif

  SingleCtx sctx;
  if (!pass) {
    if (!dom) return 1;
    sctx.rank.G = G; sctx.rank.ioff = -(G.isc - 1); sctx.rank.joff = -(G.jsc - 1);
    sctx.rank.Tr = Tr; sctx.dom = dom; sctx.ntr = ntr;
    pass = single_pass; sum = single_sum; ctx = &sctx;
  }

  const int is = G.isc, ie = G.iec, js = G.jsc, je = G.jec, nz = G.nk;       // :114
  const int isd = G.isd, ied = G.ied, jsd = G.jsd, jed = G.jed;              // :115
  const int IsdB = G.IsdB, IedB = G.IedB, JsdB = G.JsdB, JedB = G.JedB;     // :116
  int stencil = 2;                                                           // :118

  const size_t n_h = (size_t)(ied - isd + 1) * (jed - jsd + 1);
  const size_t n_u = (size_t)(IedB - IsdB + 1) * (jed - jsd + 1);
  const size_t n_v = (size_t)(ied - isd + 1) * (JedB - JsdB + 1);
  std::vector<double> hprev_(n_h * nz), uhr_(n_u * nz), vhr_(n_v * nz);
  std::vector<double> uh_neglect_(n_u, 0.0), vh_neglect_(n_v, 0.0);
  V3 hprev(hprev_.data(), isd, ied, jsd, jed), uhr(uhr_.data(), IsdB, IedB, jsd, jed), vhr(vhr_.data(), isd, ied, JsdB, JedB);
  V3 h_end(h_end_, isd, ied, jsd, jed), uhtr(uhtr_, IsdB, IedB, jsd, jed), vhtr(vhtr_, isd, ied, JsdB, JedB);
  V3 vol_prev(vol_prev_, isd, ied, jsd, jed);
  V2 uh_neglect(uh_neglect_.data(), IsdB, IedB, jsd), vh_neglect(vh_neglect_.data(), isd, ied, JsdB);
  Flags2 domore_u(jsd, jed, nz), domore_v(JsdB, JedB, nz);                   // :112-113
  std::vector<int> domore_k(nz + 1, 0);

  bool x_first = ((G.first_direction % 2) == 0);                             // :126
  const bool use_PPM_stencil = CS->usePPM && !CS->useHuynhStencilBug;        // :129
  if (use_PPM_stencil) stencil = 3;                                          // :130
  const double Idt = 1.0 / dt;                                               // :133
  int max_iter = 2 * (int)std::ceil(dt / CS->dt) + 1;                        // :135
  if (max_iter_in >= 0) max_iter = max_iter_in;                              // :137
  if (x_first_in >= 0) x_first = (x_first_in != 0);                          // :138
  const bool usePPM = CS->usePPM != 0, useHuynh = CS->useHuynh != 0;

  // :147-178
#pragma omp parallel for schedule(static)
  for (int k = 1; k <= nz; k++) {
    for (int j = jsd; j <= jed; j++) for (int I = IsdB; I <= IedB; I++) uhr(I, j, k) = 0.0;
    for (int J = JsdB; J <= JedB; J++) for (int i = isd; i <= ied; i++) vhr(i, J, k) = 0.0;
    for (int j = jsd; j <= jed; j++) for (int i = isd; i <= ied; i++) hprev(i, j, k) = 0.0;
    domore_k[k] = 1;
    for (int j = js; j <= je; j++) for (int I = is - 1; I <= ie; I++) uhr(I, j, k) = uhtr(I, j, k);
    for (int J = js - 1; J <= je; J++) for (int i = is; i <= ie; i++) vhr(i, J, k) = vhtr(i, J, k);
    if (!vol_prev_) {
      for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++) {
        hprev(i, j, k) = fmax2(0.0, C.areaT(i, j) * h_end(i, j, k) +
                                        ((uhr(i, j, k) - uhr(i - 1, j, k)) + (vhr(i, j, k) - vhr(i, j - 1, k))));
        hprev(i, j, k) = hprev(i, j, k) + fmax2(0.0, 1.0e-13 * hprev(i, j, k) - C.areaT(i, j) * h_end(i, j, k));
      }
    } else {
      for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++) hprev(i, j, k) = vol_prev(i, j, k);
    }
  }

  // :182-188
  for (int j = jsd; j <= jed; j++) for (int i = isd; i <= ied - 1; i++)
    uh_neglect(i, j) = G.H_subroundoff * fmin2(C.areaT(i, j), C.areaT(i + 1, j));
  for (int j = jsd; j <= jed - 1; j++) for (int i = isd; i <= ied; i++)
    vh_neglect(i, j) = G.H_subroundoff * fmin2(C.areaT(i, j), C.areaT(i, j + 1));

  // :192-198
  for (int m = 0; m < ntr; m++) {
    if (Tr[m].ad_x) std::memset(Tr[m].ad_x, 0, sizeof(double) * n_u * nz);
    if (Tr[m].ad_y) std::memset(Tr[m].ad_y, 0, sizeof(double) * n_v * nz);
    if (Tr[m].advection_xy) std::memset(Tr[m].advection_xy, 0, sizeof(double) * n_h * nz);
    if (Tr[m].ad2d_x) std::memset(Tr[m].ad2d_x, 0, sizeof(double) * n_u);
    if (Tr[m].ad2d_y) std::memset(Tr[m].ad2d_y, 0, sizeof(double) * n_v);
  }

  int isv = is, iev = ie, jsv = js, jev = je;                                // :201
  int itt;
  for (itt = 1; itt <= max_iter; itt++) {                                    // :203
    if (isv > is - stencil) {                                                // :205
      pass(ctx, uhr_.data(), vhr_.data(), hprev_.data(), Tr, ntr);           // :206
      int nsten_halo = std::min(std::min(is - isd, ied - ie), std::min(js - jsd, jed - je)) / stencil;   // :208
      isv = is - nsten_halo * stencil; jsv = js - nsten_halo * stencil;      // :209
      iev = ie + nsten_halo * stencil; jev = je + nsten_halo * stencil;      // :210
      if ((nsten_halo > 1) || (itt == 1)) {                                  // :213
#pragma omp parallel for schedule(static)
        for (int k = 1; k <= nz; k++) if (domore_k[k] > 0) {                 // :215
          for (int j = jsv; j <= jev; j++) if (!domore_u(j, k)) {            // :216
            for (int i = isv + stencil - 1; i <= iev - stencil; i++) if (uhr(i, j, k) != 0.0) {
              domore_u(j, k) = 1; break;
            }
          }
          for (int J = jsv + stencil - 1; J <= jev - stencil; J++) if (!domore_v(J, k)) {   // :221
            for (int i = isv + stencil; i <= iev - stencil; i++) if (vhr(i, J, k) != 0.0) {
              domore_v(J, k) = 1; break;
            }
          }
          domore_k[k] = 0;                                                   // :229
          for (int j = jsv; j <= jev; j++) if (domore_u(j, k)) domore_k[k] = 1;
          for (int J = jsv + stencil - 1; J <= jev - stencil; J++) if (domore_v(J, k)) domore_k[k] = 1;
        }
      }
    }

    isv = isv + stencil; iev = iev - stencil;                                // :238
    jsv = jsv + stencil; jev = jev - stencil;                                // :239

    if (x_first) {                                                           // :249
#pragma omp parallel for ordered schedule(static, 1)
      for (int k = 1; k <= nz; k++) if (domore_k[k] > 0) {
        advect_x(C, Tr, ntr, hprev, uhr, uh_neglect, domore_u, Idt, isv, iev, jsv - stencil, jev + stencil, k, usePPM, useHuynh);
      }
#pragma omp parallel for ordered schedule(static, 1)
      for (int k = 1; k <= nz; k++) if (domore_k[k] > 0) {
        advect_y(C, Tr, ntr, hprev, vhr, vh_neglect, domore_v, Idt, isv, iev, jsv, jev, k, usePPM, useHuynh);
        domore_k[k] = 0;                                                     // :265-267
        for (int j = jsv - stencil; j <= jev + stencil; j++) if (domore_u(j, k)) domore_k[k] = 1;
        for (int J = jsv - 1; J <= jev; J++) if (domore_v(J, k)) domore_k[k] = 1;
      }
    } else {                                                                 // :271
#pragma omp parallel for ordered schedule(static, 1)
      for (int k = 1; k <= nz; k++) if (domore_k[k] > 0) {
        advect_y(C, Tr, ntr, hprev, vhr, vh_neglect, domore_v, Idt, isv - stencil, iev + stencil, jsv, jev, k, usePPM, useHuynh);
      }
#pragma omp parallel for ordered schedule(static, 1)
      for (int k = 1; k <= nz; k++) if (domore_k[k] > 0) {
        advect_x(C, Tr, ntr, hprev, uhr, uh_neglect, domore_u, Idt, isv, iev, jsv, jev, k, usePPM, useHuynh);
        domore_k[k] = 0;                                                     // :287-289
        for (int j = jsv; j <= jev; j++) if (domore_u(j, k)) domore_k[k] = 1;
        for (int J = jsv - 1; J <= jev; J++) if (domore_v(J, k)) domore_k[k] = 1;
      }
    }

    if (itt >= max_iter) break;                                              // :297-299

    if (isv > is - stencil) {                                                // :302
      int do_any = 0;
      sum(ctx, domore_k.data() + 1, nz);                                     // :305
      for (int k = 1; k <= nz; k++) do_any = do_any + domore_k[k];
      if (do_any == 0) break;                                                // :308-310
    }
  }
  if (n_iter_out) *n_iter_out = std::min(itt, max_iter);

  if (uhr_out) std::memcpy(uhr_out, uhr_.data(), sizeof(double) * n_u * nz);             // :316
  if (vhr_out) std::memcpy(vhr_out, vhr_.data(), sizeof(double) * n_v * nz);             // :317
  if (vol_prev_ && update_vol_prev >= 0) {                                                // :318-320
    if (update_vol_prev) std::memcpy(vol_prev_, hprev_.data(), sizeof(double) * n_h * nz);
  }
  return 0;
}

extern "C" void ora_group_pass(int nranks, ora_rank *ranks, const ora_domain *dom,
                               double **uhr, double **vhr, double **hprev, int ntr) {
  // sequential emulation: gather every rank's receive values first, then write
  struct Pending { double *p; double v; };
  std::vector<Pending> pend;
  for (int r = 0; r < nranks; r++) {
    const int nk = ranks[r].G.nk;
    for (int p = 0; p < 3; p++) {
      Pos pos = (Pos)p;
      std::vector<HaloItem> list;
      build_halo_list(nranks, ranks, dom, r, pos, list);
      const long psz = plane_size(ranks[r].G, pos);
      auto gather = [&](double *dst, double **srcs, bool vec) {
        for (int k = 0; k < nk; k++) for (const HaloItem &it : list) {
          const long spsz = plane_size(ranks[it.srank].G, pos);
          double v = srcs[it.srank][it.src + spsz * k];
          pend.push_back({dst + it.dst + psz * k, (vec && it.sgn < 0.0) ? -v : v});
        }
      };
      if (pos == POS_U) gather(uhr[r], uhr, true);
      else if (pos == POS_V) gather(vhr[r], vhr, true);
      else {
        gather(hprev[r], hprev, false);
        std::vector<double *> tp(nranks);
        for (int m = 0; m < ntr; m++) {
          for (int q = 0; q < nranks; q++) tp[q] = ranks[q].Tr[m].t;
          gather(ranks[r].Tr[m].t, tp.data(), false);
        }
      }
    }
  }
  for (const Pending &q : pend) *q.p = q.v;
}

extern "C" int ora_advect_tracer_multi(int nranks, ora_rank *ranks, const ora_domain *dom,
                                       const ora_cs *CS, int ntr, double dt, int x_first_in,
                                       int max_iter_in, int update_vol_prev) {
  MultiShared sh(nranks);
  sh.nranks = nranks; sh.ranks = ranks; sh.dom = dom; sh.ntr = ntr;
  std::vector<MultiCtx> ctxs(nranks);
  std::vector<int> rc(nranks, 0);
  std::vector<std::thread> th;
  for (int r = 0; r < nranks; r++) {
    ctxs[r].sh = &sh; ctxs[r].r = r;
    th.emplace_back([&, r] {
      rc[r] = ora_advect_tracer(&ranks[r].G, CS, ranks[r].Tr, ntr, ranks[r].h_end, ranks[r].uhtr, ranks[r].vhtr, dt,
                                x_first_in, ranks[r].vol_prev, max_iter_in, update_vol_prev, ranks[r].uhr_out,
                                ranks[r].vhr_out, nullptr, multi_pass, multi_sum, &ctxs[r], &ranks[r].n_iter, 1);
    });
  }
  for (auto &t : th) t.join();
  for (int r = 0; r < nranks; r++) if (rc[r]) return rc[r];
  return 0;
}

// The same with an ocean_OBC_type per rank (NULL entries == unassociated): what MOM6 runs with open boundaries on more
// than one PE, every PE holding its own clipped, locally indexed segments (MOM_open_boundary.F90:1296-1328).
extern "C" int ora_advect_tracer_multi_obc(int nranks, ora_rank *ranks, const ora_domain *dom, const ora_cs *CS, int ntr,
                                           double dt, int x_first_in, int max_iter_in, int update_vol_prev,
                                           const ora_obc *const *obcs) {
  MultiShared sh(nranks);
  sh.nranks = nranks; sh.ranks = ranks; sh.dom = dom; sh.ntr = ntr;
  std::vector<MultiCtx> ctxs(nranks);
  std::vector<int> rc(nranks, 0);
  std::vector<std::thread> th;
  for (int r = 0; r < nranks; r++) {
    ctxs[r].sh = &sh; ctxs[r].r = r;
    th.emplace_back([&, r] {
      rc[r] = ora_advect_tracer_obc(&ranks[r].G, CS, ranks[r].Tr, ntr, ranks[r].h_end, ranks[r].uhtr, ranks[r].vhtr, dt,
                                    x_first_in, ranks[r].vol_prev, max_iter_in, update_vol_prev, ranks[r].uhr_out,
                                    ranks[r].vhr_out, nullptr, multi_pass, multi_sum, &ctxs[r], &ranks[r].n_iter, 1,
                                    obcs ? obcs[r] : nullptr);
    });
  }
  for (auto &t : th) t.join();
  for (int r = 0; r < nranks; r++) if (rc[r]) return rc[r];
  return 0;
}

// The same lock-step run with the per-rank routine supplied by the caller: `impl` has the argument list of
// ref_advect_tracer_obc (oracle/f2c/ref_wrap.cpp), i.e. the UNMODIFIED reference file translated mechanically.  This is
// how a layout question is put to the reference's own statements rather than to the restatement above.
typedef int (*ora_rank_impl_fn)(const ora_grid *, const ora_cs *, ora_tracer *, int, const double *, const double *,
                                const double *, double, int, double *, int, int, double *, double *, const ora_domain *,
                                ora_pass_fn, ora_sum_fn, void *, int *, const ora_obc *);
extern "C" int ora_advect_tracer_multi_impl(int nranks, ora_rank *ranks, const ora_domain *dom, const ora_cs *CS, int ntr,
                                            double dt, int x_first_in, int max_iter_in, int update_vol_prev,
                                            const ora_obc *const *obcs, void *impl_ptr) {
  ora_rank_impl_fn impl = (ora_rank_impl_fn)impl_ptr;
  MultiShared sh(nranks);
  sh.nranks = nranks; sh.ranks = ranks; sh.dom = dom; sh.ntr = ntr;
  std::vector<MultiCtx> ctxs(nranks);
  std::vector<int> rc(nranks, 0);
  std::vector<std::thread> th;
  for (int r = 0; r < nranks; r++) {
    ctxs[r].sh = &sh; ctxs[r].r = r;
    th.emplace_back([&, r] {
      int npass = 0;
      rc[r] = impl(&ranks[r].G, CS, ranks[r].Tr, ntr, ranks[r].h_end, ranks[r].uhtr, ranks[r].vhtr, dt, x_first_in,
                   ranks[r].vol_prev, max_iter_in, update_vol_prev, ranks[r].uhr_out, ranks[r].vhr_out, nullptr, multi_pass,
                   multi_sum, &ctxs[r], &npass, obcs ? obcs[r] : nullptr);
      ranks[r].n_iter = npass;      // (the translated routine reports its group passes, not its trips)
    });
  }
  for (auto &t : th) t.join();
  for (int r = 0; r < nranks; r++) if (rc[r]) return rc[r];
  return 0;
}

// ---------------------------------------------------------------------------
// Parity / conservation metrics restated from the reference framework.
// ---------------------------------------------------------------------------
extern "C" long long ora_bitcount_chksum(const double *a, long long n) {
  // MOM_checksums.F90:2407-2414 bitcount(abs(x)) summed, :1396-1404, mod bc_modulus (:95)
  long long s = 0;
  for (long long q = 0; q < n; q++) {
    double x = std::fabs(a[q]);
    uint64_t u; std::memcpy(&u, &x, 8);
    s += __builtin_popcountll(u);
  }
  return s % 1000000000LL;
}

extern "C" void ora_efp_sum(const double *a, long long n, long long limbs[6]) {
  // MOM_coms.F90:28-37 (prec=2**46, ni=6), real_to_ints :508-538, increment_ints :560-590
  const double r_prec = 70368744177664.0;  // 2**46
  const long long prec = 1LL << 46;
  const double pr[6] = {r_prec * r_prec, r_prec, 1.0, 1.0 / r_prec, 1.0 / (r_prec * r_prec), 1.0 / (r_prec * r_prec * r_prec)};
  const double I_pr[6] = {1.0 / (r_prec * r_prec), 1.0 / r_prec, 1.0, r_prec, r_prec * r_prec, r_prec * r_prec * r_prec};
  for (int i = 0; i < 6; i++) limbs[i] = 0;
  for (long long q = 0; q < n; q++) {
    double r = a[q];
    long long ints[6];
    int sgn = (r < 0.0) ? -1 : 1;
    double rs = std::fabs(r);
    for (int i = 0; i < 6; i++) {
      long long ival = (long long)(rs * I_pr[i]);
      rs = rs - ival * pr[i];
      ints[i] = sgn * ival;
    }
    for (int i = 5; i >= 1; i--) {
      limbs[i] += ints[i];
      if (limbs[i] > prec) { limbs[i] -= prec; limbs[i - 1] += 1; }
      else if (limbs[i] < -prec) { limbs[i] += prec; limbs[i - 1] -= 1; }
    }
    limbs[0] += ints[0];
  }
}

extern "C" double ora_efp_to_real(const long long limbs[6]) {
  const double r_prec = 70368744177664.0;
  const double pr[6] = {r_prec * r_prec, r_prec, 1.0, 1.0 / r_prec, 1.0 / (r_prec * r_prec), 1.0 / (r_prec * r_prec * r_prec)};
  double r = 0.0;
  for (int i = 0; i < 6; i++) r = r + pr[i] * (double)limbs[i];
  return r;
}

// ---------------------------------------------------------------------------
// tracer_hordiff, the along-layer diffusion iteration:
// /root/reference/src/tracer/MOM_tracer_hor_diff.F90:539-607 restated.  khdt_x / khdt_y (2-D, :186-366), num_itts
// (:368-387), h_neglect (:205) and Idt (:204) are the caller's, exactly as the loop finds them in the reference.
// Pinned against the translated reference lines (oracle/f2c, tests/test_hordiff.py).
// ---------------------------------------------------------------------------
static int hordiff_impl(const ora_grid *Gp, ora_tracer *Tr, int ntr, const double *h, const double *khdt_x,
                        const double *khdt_y, int num_itts, double h_neglect, int Diffuse_ML_interior,
                        int nkml, int nk_rho_varies, double ML_KhTr_scale, const ora_domain *dom,
                        ora_pass_fn pass, void *ctx, double Idt, double *const *df_x, double *const *df_y,
                        double *const *df2d_x, double *const *df2d_y) {
  const ora_grid &G = *Gp;
  if (!dom && !pass) return 1;
  const int is = G.isc, ie = G.iec, js = G.jsc, je = G.jec, nz = G.nk;
  const long nih = G.ied - G.isd + 1, njh = G.jed - G.jsd + 1;
  const long niu = G.IedB - G.IsdB + 1, njv = G.JedB - G.JsdB + 1;
  auto H = [&](int i, int j, int k) -> double { return h[(i - G.isd) + nih * ((j - G.jsd) + njh * (long)(k - 1))]; };
  auto KX = [&](int I, int j) -> double { return khdt_x[(I - G.IsdB) + niu * (long)(j - G.jsd)]; };
  auto KY = [&](int i, int J) -> double { return khdt_y[(i - G.isd) + nih * (long)(J - G.JsdB)]; };
  auto IA = [&](int i, int j) -> double { return G.IareaT[(i - G.isd) + nih * (long)(j - G.jsd)]; };
  // the tracer-only group pass (CS%pass_t, :540): the emulation moves uhr/vhr/hprev too -- give it scratch fields
  std::vector<double> su((size_t)niu * njh * nz, 0.0), sv((size_t)nih * njv * nz, 0.0), sh((size_t)nih * njh * nz, 0.0);
  SingleCtx sctx;
  if (!pass) {
    sctx.rank.G = G; sctx.rank.ioff = -(G.isc - 1); sctx.rank.joff = -(G.jsc - 1);
    sctx.rank.Tr = Tr; sctx.dom = dom; sctx.ntr = ntr;
    pass = single_pass; ctx = &sctx;
  }
  std::vector<double> Coef_x((size_t)niu * njh), Coef_y((size_t)nih * njv), Ihdxdy((size_t)nih * njh), dTr((size_t)nih * njh);
  auto CX = [&](int I, int j) -> double & { return Coef_x[(I - G.IsdB) + niu * (long)(j - G.jsd)]; };
  auto CY = [&](int i, int J) -> double & { return Coef_y[(i - G.isd) + nih * (long)(J - G.JsdB)]; };
  const double I_numitts = 1.0 / ((double)num_itts);                                   // :380 / :384 / :386
  for (int itt = 1; itt <= num_itts; itt++) {                                         // :539
    pass(ctx, su.data(), sv.data(), sh.data(), Tr, ntr);                              // :540
    for (int k = 1; k <= nz; k++) {                                                   // :542
      double scale = I_numitts;                                                       // :543
      if (Diffuse_ML_interior) {                                                      // :544-550
        if (k <= nkml) {
          if (ML_KhTr_scale <= 0.0) continue;
          scale = I_numitts * ML_KhTr_scale;
        }
        if ((k > nkml) && (k <= nk_rho_varies)) continue;
      }
      for (int J = js - 1; J <= je; J++) for (int i = is; i <= ie; i++)               // :552-555
        CY(i, J) = ((scale * KY(i, J)) * 2.0 * (H(i, J, k) * H(i, J + 1, k))) / (H(i, J, k) + H(i, J + 1, k) + h_neglect);
      for (int j = js; j <= je; j++) {                                                // :557-566
        for (int I = is - 1; I <= ie; I++)
          CX(I, j) = ((scale * KX(I, j)) * 2.0 * (H(I, j, k) * H(I + 1, j, k))) / (H(I, j, k) + H(I + 1, j, k) + h_neglect);
        for (int i = is; i <= ie; i++) Ihdxdy[(i - G.isd) + nih * (long)(j - G.jsd)] = IA(i, j) / (H(i, j, k) + h_neglect);
      }
      for (int m = 0; m < ntr; m++) {                                                 // :568-595
        double *t = Tr[m].t + nih * njh * (long)(k - 1);
        auto T = [&](int i, int j) -> double & { return t[(i - G.isd) + nih * (long)(j - G.jsd)]; };
        for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++)
          dTr[(i - G.isd) + nih * (long)(j - G.jsd)] = Ihdxdy[(i - G.isd) + nih * (long)(j - G.jsd)] *
              ((CX(i - 1, j) * (T(i - 1, j) - T(i, j)) - CX(i, j) * (T(i, j) - T(i + 1, j))) +
               (CY(i, j - 1) * (T(i, j - 1) - T(i, j)) - CY(i, j) * (T(i, j) - T(i, j + 1))));
        // diffusive-flux diagnostics, :576-591 (accumulated over iterations; the caller zeroes them, :388-405).
        // G%IscB = isc-1 in symmetric memory, isc otherwise (MOM_hor_index.F90:87-96); G%IecB = iec.
        {
          const int IscB = (G.IsdB == G.isd - 1) ? G.isc - 1 : G.isc, JscB = (G.JsdB == G.jsd - 1) ? G.jsc - 1 : G.jsc;
          if (df_x && df_x[m]) for (int j = js; j <= je; j++) for (int I = IscB; I <= G.iec; I++) {
            double &d = df_x[m][(I - G.IsdB) + niu * ((j - G.jsd) + njh * (long)(k - 1))];
            d = d + CX(I, j) * (T(I, j) - T(I + 1, j)) * Idt;                                      // :577-578
          }
          if (df_y && df_y[m]) for (int J = JscB; J <= G.jec; J++) for (int i = is; i <= ie; i++) {
            double &d = df_y[m][(i - G.isd) + nih * ((J - G.JsdB) + njv * (long)(k - 1))];
            d = d + CY(i, J) * (T(i, J) - T(i, J + 1)) * Idt;                                      // :581-582
          }
          if (df2d_x && df2d_x[m]) for (int j = js; j <= je; j++) for (int I = IscB; I <= G.iec; I++) {
            double &d = df2d_x[m][(I - G.IsdB) + niu * (long)(j - G.jsd)];
            d = d + CX(I, j) * (T(I, j) - T(I + 1, j)) * Idt;                                      // :585-586
          }
          if (df2d_y && df2d_y[m]) for (int J = JscB; J <= G.jec; J++) for (int i = is; i <= ie; i++) {
            double &d = df2d_y[m][(i - G.isd) + nih * (long)(J - G.JsdB)];
            d = d + CY(i, J) * (T(i, J) - T(i, J + 1)) * Idt;                                      // :589-590
          }
        }
        for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++) T(i, j) = T(i, j) + dTr[(i - G.isd) + nih * (long)(j - G.jsd)];
      }
    }
    for (int m = 0; m < ntr; m++) if (Tr[m].conc_underflow > 0.0) {                   // :600-605
      for (int k = 1; k <= nz; k++) for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++) {
        double &v = Tr[m].t[(i - G.isd) + nih * ((j - G.jsd) + njh * (long)(k - 1))];
        if (std::fabs(v) < Tr[m].conc_underflow) v = 0.0;
      }
    }
  }
  return 0;
}

extern "C" int ora_tracer_hordiff(const ora_grid *Gp, ora_tracer *Tr, int ntr, const double *h, const double *khdt_x,
                                  const double *khdt_y, int num_itts, double h_neglect, int Diffuse_ML_interior,
                                  int nkml, int nk_rho_varies, double ML_KhTr_scale, const ora_domain *dom,
                                  ora_pass_fn pass, void *ctx) {
  return hordiff_impl(Gp, Tr, ntr, h, khdt_x, khdt_y, num_itts, h_neglect, Diffuse_ML_interior, nkml, nk_rho_varies,
                      ML_KhTr_scale, dom, pass, ctx, 0.0, nullptr, nullptr, nullptr, nullptr);
}

// The same with the diffusive-flux diagnostics df_x / df_y / df2d_x / df2d_y (:576-591): arrays of ntr pointers, NULL
// entries (or a NULL array) == unassociated; Idt = 1/dt (:204).  The arrays accumulate: zero them first (:388-405).
extern "C" int ora_tracer_hordiff_df(const ora_grid *Gp, ora_tracer *Tr, int ntr, const double *h, const double *khdt_x,
                                     const double *khdt_y, int num_itts, double h_neglect, int Diffuse_ML_interior,
                                     int nkml, int nk_rho_varies, double ML_KhTr_scale, const ora_domain *dom,
                                     ora_pass_fn pass, void *ctx, double Idt, double *const *df_x, double *const *df_y,
                                     double *const *df2d_x, double *const *df2d_y) {
  return hordiff_impl(Gp, Tr, ntr, h, khdt_x, khdt_y, num_itts, h_neglect, Diffuse_ML_interior, nkml, nk_rho_varies,
                      ML_KhTr_scale, dom, pass, ctx, Idt, df_x, df_y, df2d_x, df2d_y);
}

// ---------------------------------------------------------------------------
// The iteration loop of offline_advection_ale:
// /root/reference/src/tracer/MOM_offline_main.F90:254-312 (set-up), :320-411 (loop), :414-415 (epilogue), with
// remaining_transport_sum (:615-650) and reproducing_sum_2d (src/framework/MOM_coms.F90:119-213,219-318) restated.
// The ALE step inside the loop (:346-380, other modules) is NOT restated: the call is refused unless
// MODULO(iter, off_ale_mod) /= 0 for every iteration it can run.  The debug/print branches (:327-332, :392-395) only
// write messages.  pass_var / pass_vector are the oracle's group-pass emulation on (h_new) and (uhtr_sub, vhtr_sub).
// ---------------------------------------------------------------------------
namespace {
// reproducing_sum_2d of a non-NaN array section, MOM_coms.F90:119-213 (overflow_check default .true.) + :285
double repro_sum_2d(const double *a, long nrow_stride, int i0, int i1, int j0, int j1) {
  const double r_prec = 70368744177664.0;            // 2.0**46, :28
  const long long prec = 1LL << 46;                  // :27
  const double I_prec = 1.0 / r_prec;                // :29
  const int max_count_prec = (1 << (63 - 46)) - 1;   // :30
  const double pr[6] = {r_prec * r_prec, r_prec, 1.0, 1.0 / r_prec, 1.0 / (r_prec * r_prec), 1.0 / (r_prec * r_prec * r_prec)};
  const double I_pr[6] = {1.0 / (r_prec * r_prec), 1.0 / r_prec, 1.0, r_prec, r_prec * r_prec, r_prec * r_prec * r_prec};
  long long s[6] = {0, 0, 0, 0, 0, 0};
  auto incr = [&](double r) {                        // increment_ints_faster :597-626
    const int sgn = (r < 0.0) ? -1 : 1;
    double rs = std::fabs(r);
    for (int i = 0; i < 6; i++) {
      const long long ival = (long long)(rs * I_pr[i]);
      rs = rs - ival * pr[i];
      s[i] = s[i] + sgn * ival;
    }
  };
  auto carry = [&]() {                               // carry_overflow :629-649 (the same first loop as regularize_ints)
    for (int i = 5; i >= 1; i--) if (std::llabs(s[i]) >= prec) {
      const int num_carry = (int)(s[i] * I_prec);
      s[i] = s[i] - num_carry * prec;
      s[i - 1] = s[i - 1] + num_carry;
    }
  };
  const long ni = i1 - i0 + 1, nj = j1 - j0 + 1;
  if (ni * nj < max_count_prec) {                    // :156-160
    for (int j = j0; j <= j1; j++) for (int i = i0; i <= i1; i++) incr(a[i + nrow_stride * j]);
    carry();
  } else {                                           // :161-167 (rows shorter than max_count_prec)
    for (int j = j0; j <= j1; j++) { for (int i = i0; i <= i1; i++) incr(a[i + nrow_stride * j]); carry(); }
  }
  // (sum_across_PEs of the integers: one PE) ; regularize_ints :653-691
  carry();
  bool positive = true;
  for (int i = 0; i < 6; i++) if (std::llabs(s[i]) > 0) { if (s[i] < 0) positive = false; break; }
  if (positive) { for (int i = 5; i >= 1; i--) if (s[i] < 0) { s[i] += prec; s[i - 1] -= 1; } }
  else { for (int i = 5; i >= 1; i--) if (s[i] > 0) { s[i] -= prec; s[i - 1] += 1; } }
  double r = 0.0;                                    // ints_to_real :550-559
  for (int i = 0; i < 6; i++) r = r + pr[i] * (double)s[i];
  return r;
}
}  // namespace

extern "C" double ora_remaining_transport_sum(const ora_grid *Gp, const double *uhtr, const double *vhtr, const double *h_new,
                                              double H_to_kg_m2, double L_to_m, double kg_m2_to_H, double m_to_L) {
  const ora_grid &G = *Gp;                           // MOM_offline_main.F90:615-650
  const int is = G.isc, ie = G.iec, js = G.jsc, je = G.jec, nz = G.nk;
  const long nih = G.ied - G.isd + 1, njh = G.jed - G.jsd + 1, niu = G.IedB - G.IsdB + 1, njv = G.JedB - G.JsdB + 1;
  auto U = [&](int I, int j, int k) { return uhtr[(I - G.IsdB) + niu * ((j - G.jsd) + njh * (long)(k - 1))]; };
  auto V = [&](int i, int J, int k) { return vhtr[(i - G.isd) + nih * ((J - G.JsdB) + njv * (long)(k - 1))]; };
  auto H = [&](int i, int j, int k) { return h_new[(i - G.isd) + nih * ((j - G.jsd) + njh * (long)(k - 1))]; };
  const double HL2_to_kg_scale = H_to_kg_m2 * (L_to_m * L_to_m);                            // :636
  std::vector<double> col((size_t)nih * njh, 0.0);                                          // :638
  for (int k = 1; k <= nz; k++) for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++) {   // :639-644
    const double trans_cell = (std::fabs(U(i - 1, j, k)) + std::fabs(U(i, j, k))) + (std::fabs(V(i, j - 1, k)) + std::fabs(V(i, j, k)));
    const double a = 1.0e-16 * H(i, j, k), b = G.H_subroundoff;
    const double mx = (b > a) ? b : a;               // MAX(a, b), gfortran's compare-select
    const long q = (i - G.isd) + nih * (long)(j - G.jsd);
    if (trans_cell > mx * G.areaT[q]) col[q] = col[q] + HL2_to_kg_scale * trans_cell;
  }
  // :647-648; array indices of reproducing_sum start at 1: is+(1-isd) .. ie+(1-isd)
  return 0.5 * kg_m2_to_H * (m_to_L * m_to_L) * repro_sum_2d(col.data(), nih, is - G.isd, ie - G.isd, js - G.jsd, je - G.jsd);
}

extern "C" int ora_offline_advection_loop(const ora_grid *Gp, const ora_cs *CS, ora_tracer *Tr, int ntr, double *h_pre,
                                          double *uhtr, double *vhtr, double dt_offline, int num_off_iter, int off_ale_mod,
                                          int x_before_y_in, double min_residual, double H_to_kg_m2, double L_to_m,
                                          double kg_m2_to_H, double m_to_L, const ora_domain *dom, const ora_obc *OBC,
                                          int *converged_out, int *n_iter_out, double *tot_residual_out, int nthreads) {
  const ora_grid &G = *Gp;
  if (!dom) return 1;
  for (int iter = 1; iter <= num_off_iter; iter++)
    if (off_ale_mod <= 0 || iter % off_ale_mod == 0) return 2;       // the ALE step (:346-380) is not restated
  const int is = G.isc, ie = G.iec, js = G.jsc, je = G.jec, nz = G.nk;
  const long nih = G.ied - G.isd + 1, njh = G.jed - G.jsd + 1, niu = G.IedB - G.IsdB + 1, njv = G.JedB - G.JsdB + 1;
  const size_t nh = (size_t)nih * njh * nz, nu = (size_t)niu * njh * nz, nv = (size_t)nih * njv * nz;
  auto qh = [&](int i, int j, int k) { return (size_t)((i - G.isd) + nih * ((j - G.jsd) + njh * (long)(k - 1))); };
  auto qu = [&](int I, int j, int k) { return (size_t)((I - G.IsdB) + niu * ((j - G.jsd) + njh * (long)(k - 1))); };
  auto qv = [&](int i, int J, int k) { return (size_t)((i - G.isd) + nih * ((J - G.JsdB) + njv * (long)(k - 1))); };
  bool x_before_y = x_before_y_in != 0;                                             // :254
  std::vector<double> h_new(nh, 0.0), h_vol(nh, 0.0), uhtr_sub(nu, 0.0), vhtr_sub(nv, 0.0);   // :270-273
  bool converged = false;                                                           // :276
  for (size_t q = 0; q < nu; q++) uhtr_sub[q] = uhtr[q];                            // :298-300 (whole data domain)
  for (size_t q = 0; q < nv; q++) vhtr_sub[q] = vhtr[q];                            // :301-303
  for (int k = 1; k <= nz; k++) for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++) h_new[qh(i, j, k)] = h_pre[qh(i, j, k)];   // :304-306
  double tot_residual = ora_remaining_transport_sum(Gp, uhtr, vhtr, h_new.data(), H_to_kg_m2, L_to_m, kg_m2_to_H, m_to_L);   // :312
  int nrec = 0;
  if (tot_residual_out) tot_residual_out[nrec] = tot_residual;
  nrec++;
  double prev_tot_residual = std::numeric_limits<double>::quiet_NaN();   // the reference leaves it unset before :409 (first iteration)
  SingleCtx sctx;
  sctx.rank.G = G; sctx.rank.ioff = -(G.isc - 1); sctx.rank.joff = -(G.jsc - 1); sctx.rank.Tr = Tr; sctx.dom = dom; sctx.ntr = 0;
  int iter;
  for (iter = 1; iter <= num_off_iter; iter++) {                                    // :320
    for (int k = 1; k <= nz; k++) for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++) {
      h_vol[qh(i, j, k)] = h_new[qh(i, j, k)] * G.areaT[(i - G.isd) + nih * (long)(j - G.jsd)];   // :323
      h_pre[qh(i, j, k)] = h_new[qh(i, j, k)];                                                    // :324
    }
    int rc = ora_advect_tracer_obc(Gp, CS, Tr, ntr, h_pre, uhtr_sub.data(), vhtr_sub.data(), dt_offline, x_before_y ? 1 : 0,
                                   h_vol.data(), 1, 1, uhtr, vhtr, dom, nullptr, nullptr, nullptr, nullptr, nthreads, OBC);   // :334-336
    if (rc) return rc;
    x_before_y = !x_before_y;                                                       // :339
    for (int k = 1; k <= nz; k++) for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++)
      h_new[qh(i, j, k)] = h_vol[qh(i, j, k)] * G.IareaT[(i - G.isd) + nih * (long)(j - G.jsd)];  // :342-344
    for (int k = 1; k <= nz; k++) for (int j = js; j <= je; j++) for (int i = is; i <= ie; i++) {     // :382-385
      uhtr_sub[qu(i, j, k)] = uhtr[qu(i, j, k)];
      vhtr_sub[qv(i, j, k)] = vhtr[qv(i, j, k)];
    }
    single_pass(&sctx, uhtr_sub.data(), vhtr_sub.data(), h_new.data(), Tr, 0);      // :386-387 pass_var(h_new) ; pass_vector(uhtr_sub, vhtr_sub)
    // Symmetric memory: :382-385 leaves the boundary faces is-1 / js-1 to pass_vector, which delivers the neighbouring
    // PE's (single PE: the periodic image's) value of the same face; a closed boundary has none and keeps its value.
    // (The group-pass emulation used for advect_tracer counts those faces as computational and does not move them.)
    if (dom->symmetric) {
      if (dom->cyclic_x) for (int k = 1; k <= nz; k++) for (int j = G.jsd; j <= G.jed; j++) uhtr_sub[qu(is - 1, j, k)] = uhtr_sub[qu(ie, j, k)];
      if (dom->cyclic_y) for (int k = 1; k <= nz; k++) for (int i = G.isd; i <= G.ied; i++) vhtr_sub[qv(i, js - 1, k)] = vhtr_sub[qv(i, je, k)];
    }
    tot_residual = ora_remaining_transport_sum(Gp, uhtr, vhtr, h_new.data(), H_to_kg_m2, L_to_m, kg_m2_to_H, m_to_L);   // :391
    if (tot_residual_out) tot_residual_out[nrec] = tot_residual;
    nrec++;
    if (tot_residual == 0.0) { converged = true; break; }                           // :397-402
    if ((tot_residual == prev_tot_residual) || (tot_residual < min_residual)) { converged = false; break; }   // :404-407
    prev_tot_residual = tot_residual;                                               // :409
  }
  for (size_t q = 0; q < nh; q++) h_pre[q] = h_new[q];                              // :414
  {                                                                                 // :415 pass_vector(uhtr, vhtr)
    std::vector<double> sh(nh, 0.0);
    single_pass(&sctx, uhtr, vhtr, sh.data(), Tr, 0);
  }
  if (converged_out) *converged_out = converged ? 1 : 0;
  if (n_iter_out) *n_iter_out = (iter > num_off_iter) ? num_off_iter : iter;       // iterations run
  return 0;
}
