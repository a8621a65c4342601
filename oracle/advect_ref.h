/* TEST INFRASTRUCTURE ONLY -- CPU oracle for MOM6 advect_tracer.
 *
 * C interface of the CPU restatement of
 *   /root/reference/src/tracer/MOM_tracer_advect.F90:52-1087
 * (advect_tracer, advect_x, advect_y).  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library;
 * the product path (mom6_b200/) never does.
 *
 * PARITY PIN: the reference ships no golden vectors for this path and no Fortran
 * compiler exists in the build container.  What pins this restatement is the
 * reference file itself: oracle/f2c/f2cpp.py translates the UNMODIFIED
 * MOM_tracer_advect.F90 statement by statement into C++ (oracle/_ref/, built
 * where /root/reference exists) and tests/test_f2c_reference.py requires this
 * restatement, the committed goldens and (on the GPU box) the CUDA path to equal
 * it bit for bit, whole arrays, all schemes / diagnostics / optional arguments /
 * open boundaries.  Not pinned by that: what a gfortran BINARY does where the
 * source leaves it open (MIN/MAX on +0/-0 ties, unassigned stack logicals) --
 * both are run both ways and change no compared output -- and the FMS group
 * pass, which is not in the reference tree (emulated here, see ora_group_pass).
 *
 * All arrays use the reference's Fortran layout (column major, i fastest):
 *   h-point  (isd:ied,   jsd:jed,   nk)
 *   u-point  (IsdB:IedB, jsd:jed,   nk)
 *   v-point  (isd:ied,   JsdB:JedB, nk)
 * with the integer bounds passed exactly as the Fortran model holds them.
 */
#ifndef MOM6_ADVECT_REF_H
#define MOM6_ADVECT_REF_H

#ifdef __cplusplus
extern "C" {
#endif

/* ocean_grid_type + verticalGrid_type members read by the path
 * (src/core/MOM_grid.F90:25-105, src/core/MOM_verticalGrid.F90:212-215). */
typedef struct {
  int isc, iec, jsc, jec;
  int isd, ied, jsd, jed;
  int IsdB, IedB, JsdB, JedB;
  int nk;
  int first_direction;
  const double *areaT;    /* (isd:ied, jsd:jed) */
  const double *IareaT;   /* (isd:ied, jsd:jed) */
  const double *mask2dT;  /* (isd:ied, jsd:jed) */
  const double *mask2dCu; /* (IsdB:IedB, jsd:jed) */
  const double *mask2dCv; /* (isd:ied, JsdB:JedB) */
  double Angstrom_H;
  double H_subroundoff;
} ora_grid;

/* tracer_advect_CS (MOM_tracer_advect.F90:30-40). */
typedef struct {
  double dt; /* baroclinic DT; only used for max_iter */
  int usePPM;
  int useHuynh;
  int useHuynhStencilBug;
} ora_cs;

/* tracer_type members read by the path (MOM_tracer_types.F90:9-119).
 * NULL pointer == unassociated. */
typedef struct {
  double *t;            /* (isd:ied, jsd:jed, nk) */
  double *ad_x;         /* (IsdB:IedB, jsd:jed, nk) */
  double *ad_y;         /* (isd:ied, JsdB:JedB, nk) */
  double *ad2d_x;       /* (IsdB:IedB, jsd:jed) */
  double *ad2d_y;       /* (isd:ied, JsdB:JedB) */
  double *advection_xy; /* (isd:ied, jsd:jed, nk) */
  double conc_underflow;
} ora_tracer;

/* ocean_OBC_type / OBC_segment_type / segment tracer registry members read by the path
 * (MOM_tracer_advect.F90:442-478,580-628,825-862,965-1014; src/core/MOM_open_boundary.F90:119-142,254,333).
 * tres has the segment's own bounds: E/W segments (IsdB:IedB, jsd:jed, nk), N/S (isd:ied, JsdB:JedB, nk);
 * NULL == not allocated.  ntr_index is 1-based as in Fortran.  ntseg < 0 == tr_Reg unassociated. */
enum { ORA_OBC_NONE = 0, ORA_OBC_DIRECTION_N = 100, ORA_OBC_DIRECTION_S = 200, ORA_OBC_DIRECTION_E = 300,
       ORA_OBC_DIRECTION_W = 400 };
typedef struct {
  int ntr_index;
  const double *tres;
  double OBC_inflow_conc;
} ora_obc_tracer;
typedef struct {
  int is_E_or_W, is_N_or_S, direction, specified;
  int isd, ied, jsd, jed, IsdB, IedB, JsdB, JedB;   /* segment%HI */
  int ntseg;
  const ora_obc_tracer *Tr;
} ora_obc_segment;
typedef struct {
  int OBC_pe;
  int number_of_segments;
  int open_u_BCs_exist_globally, open_v_BCs_exist_globally;
  int specified_u_BCs_exist_globally, specified_v_BCs_exist_globally;
  const ora_obc_segment *segment;
} ora_obc;

/* Stand-ins for do_group_pass (MOM_tracer_advect.F90:206) and
 * sum_across_PEs (:305).  ctx is passed through untouched. */
typedef void (*ora_pass_fn)(void *ctx, double *uhr, double *vhr, double *hprev,
                            ora_tracer *Tr, int ntr);
typedef void (*ora_sum_fn)(void *ctx, int *domore_k, int nz);

/* Domain description used by the built-in group-pass emulation: a pi x pj
 * block layout of a global nig x njg grid, MOM_domains.F90:184-191. */
typedef struct {
  int nig, njg;       /* global computational-domain size */
  int cyclic_x, cyclic_y, tripolar_n;
  int symmetric;      /* symmetric memory (u/v arrays carry the extra 0 index) */
} ora_domain;

/* One rank of an emulated multi-PE run. */
typedef struct {
  ora_grid G;
  int ioff, joff;     /* global index = local index + off (of h points) */
  ora_tracer *Tr;     /* ntr entries */
  const double *h_end, *uhtr, *vhtr;
  double *vol_prev;   /* nullable */
  double *uhr_out, *vhr_out; /* nullable */
  int n_iter;         /* out: iterations taken */
} ora_rank;

/* advect_tracer, MOM_tracer_advect.F90:52-324.
 *  x_first_in: -1 absent, 0 false, 1 true.   max_iter_in: <0 absent.
 *  update_vol_prev: -1 absent, 0 false, 1 true.
 *  pass/sum may be NULL only if dom != NULL (built-in single-rank pass).
 *  nthreads: OpenMP threads over k (where the reference threads, :147,214,247).
 * returns 0, or 1 if CS/Reg unassociated (the reference's FATAL). */
int ora_advect_tracer(const ora_grid *G, const ora_cs *CS, ora_tracer *Tr, int ntr,
                      const double *h_end, const double *uhtr, const double *vhtr,
                      double dt, int x_first_in, double *vol_prev, int max_iter_in,
                      int update_vol_prev, double *uhr_out, double *vhr_out,
                      const ora_domain *dom, ora_pass_fn pass, ora_sum_fn sum,
                      void *ctx, int *n_iter_out, int nthreads);

/* The same with an associated OBC (NULL == unassociated). */
int ora_advect_tracer_obc(const ora_grid *G, const ora_cs *CS, ora_tracer *Tr, int ntr,
                          const double *h_end, const double *uhtr, const double *vhtr,
                          double dt, int x_first_in, double *vol_prev, int max_iter_in,
                          int update_vol_prev, double *uhr_out, double *vhr_out,
                          const ora_domain *dom, ora_pass_fn pass, ora_sum_fn sum,
                          void *ctx, int *n_iter_out, int nthreads, const ora_obc *OBC);

/* Emulated multi-PE run: all ranks advance in lock step (one thread each),
 * exchanging halos through the emulated group pass.  Same optional-argument
 * conventions as above. */
int ora_advect_tracer_multi(int nranks, ora_rank *ranks, const ora_domain *dom,
                            const ora_cs *CS, int ntr, double dt, int x_first_in,
                            int max_iter_in, int update_vol_prev);

/* The same with one ocean_OBC_type per rank (NULL == unassociated on that rank). */
int ora_advect_tracer_multi_obc(int nranks, ora_rank *ranks, const ora_domain *dom, const ora_cs *CS, int ntr, double dt,
                                int x_first_in, int max_iter_in, int update_vol_prev, const ora_obc *const *obcs);

/* The same lock-step run with the per-rank routine supplied by the caller (impl: a function with the argument list of
 * ref_advect_tracer_obc, oracle/f2c/ref_wrap.cpp -- the mechanically translated reference file). */
int ora_advect_tracer_multi_impl(int nranks, ora_rank *ranks, const ora_domain *dom, const ora_cs *CS, int ntr, double dt,
                                 int x_first_in, int max_iter_in, int update_vol_prev, const ora_obc *const *obcs, void *impl);

/* Emulated do_group_pass for an arbitrary set of ranks (exposed for tests).
 * fields of rank r: uhr[r], vhr[r] (C-grid vector pair), hprev[r], Tr of rank. */
void ora_group_pass(int nranks, ora_rank *ranks, const ora_domain *dom,
                    double **uhr, double **vhr, double **hprev, int ntr);

/* tracer_hordiff, the along-layer diffusion iteration: MOM_tracer_hor_diff.F90:539-607.
 * h (isd:ied, jsd:jed, nk); khdt_x (IsdB:IedB, jsd:jed) and khdt_y (isd:ied, JsdB:JedB) as the reference's 2-D set-up
 * (:186-366) leaves them; num_itts (:368-387); h_neglect = GV%H_subroundoff (:205).  df_* diagnostics are not restated. */
int ora_tracer_hordiff(const ora_grid *G, ora_tracer *Tr, int ntr, const double *h, const double *khdt_x,
                       const double *khdt_y, int num_itts, double h_neglect, int Diffuse_ML_interior, int nkml,
                       int nk_rho_varies, double ML_KhTr_scale, const ora_domain *dom, ora_pass_fn pass, void *ctx);

/* The iteration loop of offline_advection_ale: MOM_offline_main.F90:254-312, 320-411, 414-415,
 * without the ALE step inside the loop (:346-380): returns 2 if MODULO(iter, off_ale_mod) == 0 for an iteration it may
 * run.  h_pre, uhtr, vhtr are the routine's intent(inout) arrays (whole data domain); tot_residual_out (nullable,
 * num_off_iter+1 values) receives remaining_transport_sum at :312 and at :391 of every iteration run. */
int ora_offline_advection_loop(const ora_grid *G, const ora_cs *CS, ora_tracer *Tr, int ntr, double *h_pre, double *uhtr,
                               double *vhtr, double dt_offline, int num_off_iter, int off_ale_mod, int x_before_y,
                               double min_residual, double H_to_kg_m2, double L_to_m, double kg_m2_to_H, double m_to_L,
                               const ora_domain *dom, const ora_obc *OBC, int *converged, int *n_iter,
                               double *tot_residual_out, int nthreads);
/* remaining_transport_sum, MOM_offline_main.F90:615-650 (reproducing_sum_2d of MOM_coms.F90 restated) */
double ora_remaining_transport_sum(const ora_grid *G, const double *uhtr, const double *vhtr, const double *h_new,
                                   double H_to_kg_m2, double L_to_m, double kg_m2_to_H, double m_to_L);

/* tracer_hordiff's iteration with its diffusive-flux diagnostics df_x / df_y / df2d_x / df2d_y (:576-591): arrays of ntr
 * pointers, NULL == unassociated; Idt = 1/dt; the arrays accumulate over iterations and must be zeroed first (:388-405). */
int ora_tracer_hordiff_df(const ora_grid *G, ora_tracer *Tr, int ntr, const double *h, const double *khdt_x,
                          const double *khdt_y, int num_itts, double h_neglect, int Diffuse_ML_interior, int nkml,
                          int nk_rho_varies, double ML_KhTr_scale, const ora_domain *dom, ora_pass_fn pass, void *ctx,
                          double Idt, double *const *df_x, double *const *df_y, double *const *df2d_x, double *const *df2d_y);

/* MOM bitcount checksum of a 2-D/3-D array section
 * (src/framework/MOM_checksums.F90:2407-2414 bitcount, :1396-1404 use). */
long long ora_bitcount_chksum(const double *a, long long n);

/* Extended-fixed-point reproducing sum (src/framework/MOM_coms.F90:28-37,
 * 508-560): returns the six int64 limbs of sum(a[0..n)). */
void ora_efp_sum(const double *a, long long n, long long limbs[6]);
double ora_efp_to_real(const long long limbs[6]);

#ifdef __cplusplus
}
#endif
#endif
