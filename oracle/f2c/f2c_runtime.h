// TEST INFRASTRUCTURE ONLY -- C++ runtime for the code oracle/f2c/f2cpp.py generates from the UNMODIFIED
// reference source (/root/reference/src/tracer/MOM_tracer_advect.F90).  It plays the role of the stand-in Fortran
// modules of oracle/fortran_stub_harness: the derived types with the members the path reads,
// Fortran-bounds arrays, optional arguments, and the intrinsics with gfortran's semantics.  Nothing here contains
// arithmetic of the path itself -- that comes, statement by statement, from the reference file.
#pragma once
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <initializer_list>
#include <limits>
#include <memory>
#include <string>
#include <utility>
#include <vector>

typedef unsigned char lbool;      // Fortran LOGICAL

// What an automatic (stack) LOGICAL array holds before it is assigned.  The reference reads do_i(is-1,j),
// do_i(ie+1,j), do_i(i,js-1), do_i(i,je+1) unassigned (MOM_tracer_advect.F90:665,694,1072,1080; SURVEY hazard 7);
// tests build the library with both fills to show which outputs depend on it.
#ifndef F2C_LOGICAL_FILL
#define F2C_LOGICAL_FILL 0
#endif
// MIN/MAX tie rule.  gfortran >= 9 lowers MIN/MAX of reals to MIN_EXPR/MAX_EXPR ("whatever is fastest",
// gcc/fortran/trans-intrinsic.c:gfc_conv_intrinsic_minmax), i.e. x86 minsd/maxsd, whose result for equal operands
// (+0 vs -0) is the SECOND source operand; older gfortran used "m = a1; if (a2 < m) m = a2" (FIRST argument wins).
// Only +0/-0 ties can tell the two apart.  0: first argument wins (what oracle/advect_ref.cpp and the CUDA code do);
// 1: the later argument wins.  tests/test_f2c_reference.py runs both.
#ifndef F2C_MINMAX_LATER_WINS
#define F2C_MINMAX_LATER_WINS 0
#endif

template <class T> inline T f_uninit();
template <> inline double f_uninit<double>() { return std::numeric_limits<double>::quiet_NaN(); }
template <> inline int f_uninit<int>() { return -987654321; }
template <> inline lbool f_uninit<lbool>() { return (lbool)F2C_LOGICAL_FILL; }

struct Bnd { int lo, hi; };

template <class T, int N>
struct Arr {
  T *d = nullptr;
  int lo[N] = {}, hi[N] = {};
  long long st[N] = {};
  std::shared_ptr<std::vector<T>> own;

  void set_bounds(const Bnd *b) {
    long long s = 1;
    for (int n = 0; n < N; n++) { lo[n] = b[n].lo; hi[n] = b[n].hi; st[n] = s; s *= (long long)std::max(0, hi[n] - lo[n] + 1); }
  }
  long long size() const { long long s = 1; for (int n = 0; n < N; n++) s *= (long long)std::max(0, hi[n] - lo[n] + 1); return s; }
  static Arr automatic(std::initializer_list<Bnd> b) {
    Arr a; a.set_bounds(b.begin());
    a.own = std::make_shared<std::vector<T>>((size_t)std::max<long long>(a.size(), 1), f_uninit_elem());
    a.d = a.own->data();
    return a;
  }
  static T f_uninit_elem() { return elem_fill((T *)nullptr); }
  template <class U> static U elem_fill(U *) { return U(); }
  static double elem_fill(double *) { return f_uninit<double>(); }
  static int elem_fill(int *) { return f_uninit<int>(); }
  static lbool elem_fill(lbool *) { return f_uninit<lbool>(); }
  // a view of existing storage with Fortran bounds (explicit-shape dummy argument); null stays null (absent / unassociated)
  static Arr view(T *p, std::initializer_list<Bnd> b) { Arr a; a.set_bounds(b.begin()); a.d = p; return a; }
  Arr rebound(std::initializer_list<Bnd> b) const { Arr a = *this; a.set_bounds(b.begin()); return a; }
  bool associated() const { return d != nullptr; }

  inline long long off(const int *ix) const {
    long long o = 0;
    for (int n = 0; n < N; n++) {
#ifdef F2C_BOUNDS
      if (ix[n] < lo[n] || ix[n] > hi[n]) { fprintf(stderr, "f2c: index %d of dim %d outside %d:%d\n", ix[n], n + 1, lo[n], hi[n]); abort(); }
#endif
      o += (ix[n] - lo[n]) * st[n];
    }
    return o;
  }
  T &operator()(int i) const { static_assert(N == 1, "rank"); const int ix[1] = {i}; return d[off(ix)]; }
  T &operator()(int i, int j) const { static_assert(N == 2, "rank"); const int ix[2] = {i, j}; return d[off(ix)]; }
  T &operator()(int i, int j, int k) const { static_assert(N == 3, "rank"); const int ix[3] = {i, j, k}; return d[off(ix)]; }
};

template <class T>
struct Opt {
  bool has = false; T v = T();
  Opt() {}
  Opt(T x) : has(true), v(x) {}
  operator T() const { return v; }
};
template <class T> inline bool f_present(const Opt<T> &o) { return o.has; }
template <class T, int N> inline bool f_present(const Arr<T, N> &a) { return a.d != nullptr; }
template <class T, int N> inline bool f_associated(const Arr<T, N> &a) { return a.d != nullptr; }
template <class T, int N> inline bool f_allocated(const Arr<T, N> &a) { return a.d != nullptr; }
template <class T> inline bool f_associated(const T *p) { return p != nullptr; }

template <class T, int N, class S> inline void f_assign(Arr<T, N> a, S v) { const long long n = a.size(); for (long long i = 0; i < n; i++) a.d[i] = (T)v; }
template <class T, int N> inline void f_assign(Arr<T, N> a, Arr<T, N> b) {
  const long long n = a.size();
  if (b.size() != n) { fprintf(stderr, "f2c: array assignment of different sizes\n"); abort(); }
  memmove(a.d, b.d, sizeof(T) * (size_t)n);
}

// ---- intrinsics ------------------------------------------------------------------------------------------------
template <class T> inline T f_max2(T m, T a) { return F2C_MINMAX_LATER_WINS ? ((m > a) ? m : a) : ((a > m) ? a : m); }
template <class T> inline T f_min2(T m, T a) { return F2C_MINMAX_LATER_WINS ? ((m < a) ? m : a) : ((a < m) ? a : m); }
inline double f_max(double a, double b) { return f_max2(a, b); }
inline double f_max(double a, double b, double c) { return f_max2(f_max2(a, b), c); }
inline double f_min(double a, double b) { return f_min2(a, b); }
inline double f_min(double a, double b, double c) { return f_min2(f_min2(a, b), c); }
inline int f_max(int a, int b) { return f_max2(a, b); }
inline int f_min(int a, int b) { return f_min2(a, b); }
inline int f_min(int a, int b, int c) { return f_min2(f_min2(a, b), c); }
inline int f_min(int a, int b, int c, int e) { return f_min2(f_min2(f_min2(a, b), c), e); }
inline int f_max(int a, int b, int c) { return f_max2(f_max2(a, b), c); }
inline double f_abs(double a) { return std::fabs(a); }
inline int f_abs(int a) { return a < 0 ? -a : a; }
inline double f_sign(double a, double b) { return std::copysign(a, b); }      // gfortran: copysign (the sign BIT of b)
inline int f_mod(int a, int b) { return a % b; }
inline int f_int(double a) { return (int)a; }
inline int f_int(int a) { return a; }
inline int f_ceiling(double a) { return (int)std::ceil(a); }
inline double f_tiny(double) { return DBL_MIN; }
inline double f_sqrt(double a) { return std::sqrt(a); }
inline double f_epsilon(double) { return DBL_EPSILON; }
inline double f_real(int a) { return (double)a; }
inline double f_real(double a) { return a; }
inline double f_pow(double a, int b) { if (b == 2) return a * a; if (b == 3) return (a * a) * a; return std::pow(a, b); }
// keyword arguments name=value of a call arrive as kw_name(value): a tag the callee's C++ overload may use; these pass through
template <class T> inline T kw_clock(T v) { return v; }
inline int f_trips(int lo, int hi, int st) { const int n = (hi - lo + st) / st; return n > 0 ? n : 0; }

// ---- stand-ins for the modules MOM_tracer_advect.F90 uses (members as it reads them, names case-folded) ------------
struct hor_index_type { int isd = 0, ied = 0, jsd = 0, jed = 0, isdb = 0, iedb = 0, jsdb = 0, jedb = 0; };
struct mom_domain_type { int dummy = 0; };
struct ocean_grid_type {
  int isc = 0, iec = 0, jsc = 0, jec = 0, isd = 0, ied = 0, jsd = 0, jed = 0, isdb = 0, iedb = 0, jsdb = 0, jedb = 0, ke = 0;
  int iscb = 0, iecb = 0, jscb = 0, jecb = 0;
  int first_direction = 0;
  int hi = 0;      // G%HI (hor_index_type): only handed to checksum calls
  mom_domain_type *domain = nullptr;
  Arr<double, 2> areat, iareat, mask2dt, mask2dcu, mask2dcv;
};
struct verticalgrid_type { int ke = 0; double angstrom_h = 0, h_subroundoff = 0; int nkml = 0, nk_rho_varies = 0;
                           double h_to_mks = 1.0, h_to_kg_m2 = 1.0, kg_m2_to_h = 1.0;       // MOM_offline_main.F90:636,647
                           double h_to_m = 1.0; };                                            // MOM_tracer_registry.F90:756
struct unit_scale_type { double s_to_t = 1.0, l_to_m = 1.0, m_to_l = 1.0; };
struct tracer_type {
  Arr<double, 3> t, ad_x, ad_y, advection_xy;
  Arr<double, 2> ad2d_x, ad2d_y;
  Arr<double, 3> df_x, df_y;          // MOM_tracer_hor_diff.F90:576-591
  Arr<double, 2> df2d_x, df2d_y;
  double conc_underflow = 0.0;
};
struct tracer_registry_type { int ntr = 0; Arr<tracer_type, 1> tr; };
struct group_pass_type {
  double *u = nullptr, *v = nullptr;
  std::vector<double *> scalars;
};
struct diag_ctrl { int dummy = 0; };
struct tracer_advect_cs {
  double dt = 0;
  diag_ctrl *diag = nullptr;
  lbool debug = 0, useppm = 0, usehuynh = 0, usehuynhstencilbug = 0;
  group_pass_type pass_uhr_vhr_t_hprev;
};
// tracer_hor_diff_CS members the along-layer loop reads (MOM_tracer_hor_diff.F90:539-607)
struct tracer_hor_diff_cs {
  lbool diffuse_ml_interior = 0, show_call_tree = 0;
  double ml_khtr_scale = 0.0;
  group_pass_type pass_t;
  lbool check_diffusive_cfl = 0;      // :368-387
  double max_diff_cfl = -1.0;
  int id_cfl = -1, diag = 0;
};
struct obc_segment_tracer_type { Arr<double, 3> tres; double obc_inflow_conc = 0; int ntr_index = 0; };
struct segment_tracer_registry_type { int ntseg = 0; Arr<obc_segment_tracer_type, 1> tr; };
struct obc_segment_type {
  segment_tracer_registry_type *tr_reg = nullptr;
  lbool is_e_or_w = 0, is_n_or_s = 0, specified = 0;
  int direction = 0;
  hor_index_type hi;
};
struct ocean_obc_type {
  lbool obc_pe = 0;
  int number_of_segments = 0;
  lbool open_u_bcs_exist_globally = 0, open_v_bcs_exist_globally = 0;
  lbool specified_u_bcs_exist_globally = 0, specified_v_bcs_exist_globally = 0;
  Arr<obc_segment_type, 1> segment;
};
// MOM_open_boundary.F90:79-83
static const int obc_none = 0, obc_direction_n = 100, obc_direction_s = 200, obc_direction_e = 300, obc_direction_w = 400;
static const int fatal = 1, warning = 2;
static int id_clock_advect = 1, id_clock_pass = 2, id_clock_sync = 3;
#define F2C_HAS_HORDIFF 1

inline void cpu_clock_begin(int) {}
inline void cpu_clock_end(int) {}
inline std::string f_concat(const std::string &a, const std::string &b) { return a + b; }
inline void f_mom_error(int level, const std::string &m) {
  const char *msg = m.c_str();
  fprintf(stderr, "f2c MOM_error(%s): %s\n", level == fatal ? "FATAL" : "WARNING", msg);
  if (level == fatal) throw std::string(msg);
}

// MOM_domains stand-in: the group records its fields (create_group_pass appends each time it is called, exactly like
// MOM_domain_infra.F90:1084-1137) and do_group_pass hands them to the host's exchange (the same emulated full-width
// CGRID_NE pass the oracle uses: halo.cu / advect_ref.cpp ora_group_pass).
struct f2c_host {
  void (*pass)(void *ctx, group_pass_type &g) = nullptr;
  void (*sum)(void *ctx, int *v, int n) = nullptr;
  void *ctx = nullptr;
  int n_pass = 0;
};
extern thread_local f2c_host *f2c_current_host;
inline void create_group_pass(group_pass_type &g, Arr<double, 3> u, Arr<double, 3> v, mom_domain_type *) { g.u = u.d; g.v = v.d; }
inline void create_group_pass(group_pass_type &g, Arr<double, 3> s, mom_domain_type *) { g.scalars.push_back(s.d); }
inline void do_group_pass(group_pass_type &g, mom_domain_type *, int /*clock*/ = 0) {
  f2c_current_host->n_pass++;
  if (f2c_current_host->pass) f2c_current_host->pass(f2c_current_host->ctx, g);
}
inline void sum_across_pes(Arr<int, 1> v, int n) { if (f2c_current_host->sum) f2c_current_host->sum(f2c_current_host->ctx, v.d, n); }
