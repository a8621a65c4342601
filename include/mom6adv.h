/* mom6adv.h -- C ABI of the B200-native MOM6 advect_tracer.
 *
 * This is the drop-in boundary.  A Fortran shim module
 * named MOM_tracer_advect (fortran/MOM_tracer_advect.F90 in this repo) keeps the
 * reference's public surface
 *     advect_tracer       /root/reference/src/tracer/MOM_tracer_advect.F90:52-86
 *     tracer_advect_init  /root/reference/src/tracer/MOM_tracer_advect.F90:1090
 *     tracer_advect_end   /root/reference/src/tracer/MOM_tracer_advect.F90:1153
 * and binds the entry points below through ISO_C_BINDING (the tree's own
 * precedent for bind(c) is src/framework/posix.F90:52-229).
 *
 * Conventions
 *  - plain C: pointers, sizes, PODs.  No torch / C++ types.
 *  - every function returns an int status (MOM6ADV_OK == 0).  The shim maps a
 *    non-zero status to MOM_error(FATAL, mom6adv_last_error(plan)).
 *  - all user arrays are in the reference's Fortran layout (column major, i
 *    fastest) with the bounds of mom6adv_grid:
 *        h-point (isd:ied, jsd:jed, nk)      u-point (IsdB:IedB, jsd:jed, nk)
 *        v-point (isd:ied, JsdB:JedB, nk)
 *    and may be HOST or DEVICE pointers (detected with
 *    cudaPointerGetAttributes).  Inside the plan every field lives in HBM as
 *    SoA in a padded, 128-byte aligned layout.
 *  - tracer indices m are 0-based here (the shim passes m-1).
 *  - there is NO CPU fallback: without a CUDA device plan_create fails.
 */
#ifndef MOM6ADV_H
#define MOM6ADV_H

#ifdef __cplusplus
extern "C" {
#endif

enum {
  MOM6ADV_OK = 0,
  MOM6ADV_ERR_NOT_INIT = 1,    /* plan NULL / tracer not bound (reference FATALs :120-123) */
  MOM6ADV_ERR_INVALID = 2,     /* bad argument */
  MOM6ADV_ERR_UNSUPPORTED = 3, /* halo < stencil; OBC on a multi-block plan */
  MOM6ADV_ERR_CUDA = 4,
  MOM6ADV_ERR_COMM = 5         /* halo-exchange transport failure */
};

typedef struct mom6adv_plan mom6adv_plan;

/* ocean_grid_type + verticalGrid_type members read by the path
 * (src/core/MOM_grid.F90:25-105; src/core/MOM_verticalGrid.F90:212-215).
 * Index values are passed exactly as the Fortran model holds them. */
typedef struct {
  int isc, iec, jsc, jec;
  int isd, ied, jsd, jed;
  int IsdB, IedB, JsdB, JedB;
  int nk;
  const double *areaT;    /* (isd:ied, jsd:jed)   host pointer */
  const double *IareaT;   /* (isd:ied, jsd:jed)   */
  const double *mask2dT;  /* (isd:ied, jsd:jed)   */
  const double *mask2dCu; /* (IsdB:IedB, jsd:jed) */
  const double *mask2dCv; /* (isd:ied, JsdB:JedB) */
  double Angstrom_H;
  double H_subroundoff;
} mom6adv_grid;

/* tracer_advect_CS (MOM_tracer_advect.F90:30-40) + G%first_direction. */
typedef struct {
  double dt_baroclinic;   /* CS%dt: only used for max_iter (:135) */
  int usePPM;
  int useHuynh;
  int useHuynhStencilBug;
  int debug;
  int first_direction;    /* G%first_direction (:126) */
} mom6adv_config;

/* MOM_domain_type as far as the group pass needs it
 * (src/framework/MOM_domains.F90:184-191; create_group_pass / do_group_pass,
 * config_src/infra/FMS2/MOM_domain_infra.F90:983-1158). */
typedef struct {
  int nig, njg;           /* global computational-domain size */
  int npi, npj;           /* LAYOUT */
  int pi, pj;             /* this rank's block */
  int ioff, joff;         /* global index = local index + off */
  int cyclic_x, cyclic_y, tripolar_n;
} mom6adv_domain;

/* ---- life cycle --------------------------------------------------------- */
/* device < 0: use the current CUDA device.  dom == NULL: a single closed
 * block.  Stands in for tracer_advect_init's allocate(CS) (:1107) plus the
 * per-call create_group_pass (:140-144), done once. */
int mom6adv_plan_create(const mom6adv_grid *g, const mom6adv_config *c,
                        const mom6adv_domain *dom, int device, mom6adv_plan **out);
int mom6adv_plan_destroy(mom6adv_plan *plan);   /* tracer_advect_end :1153 */
const char *mom6adv_last_error(const mom6adv_plan *plan);
const char *mom6adv_version(void);

/* ---- multi-GPU: LAYOUT = npi,npj, one rank per GPU ------------------------ */
/* The device-resident stand-in for MOM_domains' create_group_pass /
 * do_group_pass (MOM_tracer_advect.F90:140-144,206 -> mpp_do_group_update) and
 * sum_across_PEs(domore_k, nz) (:305 -> mpp_sum): halo strips are packed on
 * the GPU, exchanged rank-to-rank with NCCL send/recv over NVLink and unpacked
 * on the GPU; the convergence flags are an NCCL all-reduce.
 * Rank r = pi + npi*pj.  One rank calls mom6adv_comm_unique_id and distributes
 * the 128 bytes (MPI_Bcast in the Fortran model); every rank then calls
 * mom6adv_comm_init with the global compute extents of each layout column /
 * row (what mpp_get_compute_domains reports): ibeg/iend[npi], jbeg/jend[npj]. */
int mom6adv_comm_unique_id(char *id128);
int mom6adv_comm_init(mom6adv_plan *plan, const char *id128, const int *ibeg, const int *iend,
                      const int *jbeg, const int *jend);
/* Host-only view of the exchange plan of rank (dom->pi, dom->pj) (tests; needs
 * no device).  what: 0 self dst, 1 self src, 2 self sign, 3 send src to `peer`,
 * 4 send sign to `peer`, 5 recv dst from `peer`, 6 / 7 the index of those same
 * points in `peer`'s own padded plane and their sign (the peer-pull form of the
 * exchange: each halo value is loaded straight from the owner's field through a
 * CUDA IPC mapping); indices are offsets into one padded plane.  Returns the list length (fills at most cap entries); geom
 * (nullable) receives P, R, coff, roff of the padded plane. */
int mom6adv_exchange_list(const mom6adv_grid *g, const mom6adv_domain *dom, const int *ibeg, const int *iend,
                          const int *jbeg, const int *jend, int peer, int pos, int what, int *out, int cap,
                          int *geom);

/* ---- open boundaries (ocean_OBC_type, src/core/MOM_open_boundary.F90:119-142,145-260,264-400) ----
 * The members advect_x / advect_y read (MOM_tracer_advect.F90:442-478,580-628,825-862,965-1014).
 * tres has the segment's own bounds, Fortran layout: E/W segments (IsdB:IedB, jsd:jed, nk), N/S segments
 * (isd:ied, JsdB:JedB, nk); NULL == not allocated.  ntr_index is 1-based as in Fortran.  ntseg < 0 ==
 * segment%tr_Reg unassociated.  direction: OBC_DIRECTION_N/S/E/W = 100/200/300/400 (:80-83). */
typedef struct {
  int ntr_index;
  const double *tres;
  double OBC_inflow_conc;
} mom6adv_obc_tracer;
typedef struct {
  int is_E_or_W, is_N_or_S, direction, specified;
  int isd, ied, jsd, jed, IsdB, IedB, JsdB, JedB;     /* segment%HI */
  int ntseg;
  const mom6adv_obc_tracer *Tr;
} mom6adv_obc_segment;
typedef struct {
  int OBC_pe;
  int number_of_segments;
  int open_u_BCs_exist_globally, open_v_BCs_exist_globally;
  int specified_u_BCs_exist_globally, specified_v_BCs_exist_globally;
  const mom6adv_obc_segment *segment;
} mom6adv_obc;
/* The OBC argument of advect_tracer (:64): call before mom6adv_advect / mom6adv_advect_resident whenever the
 * reservoirs changed (every step in a model with evolving tres); the tables and reservoirs are copied to the
 * device.  NULL, or OBC_pe == 0, is the unassociated pointer.  With an active OBC the sweeps take the
 * open-boundary kernels (obc.cu). */
int mom6adv_set_obc(mom6adv_plan *plan, const mom6adv_obc *obc);

/* ---- registry data (MOM_tracer_types.F90:9-130) -------------------------- */
/* Declare Reg%ntr.  Tracer slots 0..ntr-1 get HBM storage. */
int mom6adv_set_ntr(mom6adv_plan *plan, int ntr);
/* Bind Reg%Tr(m): t is where MIRROR-mode calls read/write the tracer (host or
 * device pointer, Fortran layout); diagnostic pointers NULL == unassociated. */
int mom6adv_bind_tracer(mom6adv_plan *plan, int m, double *t, double conc_underflow,
                        double *ad_x, double *ad_y, double *ad2d_x, double *ad2d_y,
                        double *advection_xy);

/* ---- MIRROR mode: one call == one reference advect_tracer call ----------- */
/* Copies h_end/uhtr/vhtr(/vol_prev) and every bound tracer to HBM, runs the
 * whole advect_tracer algorithm (:112-320) on the GPU and copies tracers,
 * associated diagnostics and the optional outputs back.
 *   x_first: -1 absent (use first_direction), 0/1.   max_iter_in: <0 absent.
 *   vol_prev NULL == absent;  update_vol_prev: -1 absent, 0/1.
 *   uhr_out / vhr_out NULL == absent.   n_iter (nullable): iterations done. */
int mom6adv_advect(mom6adv_plan *plan, const double *h_end, const double *uhtr,
                   const double *vhtr, double dt, int x_first, double *vol_prev,
                   int max_iter_in, int update_vol_prev, double *uhr_out,
                   double *vhr_out, int *n_iter);

/* ---- RESIDENT mode: fields stay in HBM between calls --------------------- */
enum { MOM6ADV_F_H_END = 0, MOM6ADV_F_UHTR = 1, MOM6ADV_F_VHTR = 2, MOM6ADV_F_VOL_PREV = 3,
       MOM6ADV_F_UHR = 4, MOM6ADV_F_VHR = 5, MOM6ADV_F_HPREV = 6 };
int mom6adv_upload_tracer(mom6adv_plan *plan, int m, const double *src);
int mom6adv_download_tracer(mom6adv_plan *plan, int m, double *dst);
int mom6adv_upload_field(mom6adv_plan *plan, int field, const double *src);
int mom6adv_download_field(mom6adv_plan *plan, int field, double *dst);
/* kind: 0 ad_x, 1 ad_y, 2 ad2d_x, 3 ad2d_y, 4 advection_xy; 5 df_x, 6 df_y, 7 df2d_x, 8 df2d_y (mom6adv_bind_tracer_df) */
int mom6adv_download_diag(mom6adv_plan *plan, int m, int kind, double *dst);
/* Runs the algorithm on the resident fields; no host<->device traffic except
 * the per-iteration convergence word.  use_vol_prev: take hprev from the
 * VOL_PREV field (:174-176) instead of reconstructing it (:164-172). */
int mom6adv_advect_resident(mom6adv_plan *plan, double dt, int x_first, int max_iter_in,
                            int use_vol_prev, int update_vol_prev, int *n_iter);
int mom6adv_sync(mom6adv_plan *plan);

/* ---- either side of advect_tracer: what keeps its inputs and outputs off PCIe ------------
 * mom6adv_accumulate_transports replaces the loop of step_MOM_dyn_split_RK2
 *   uhtr(I,j,k) = uhtr(I,j,k) + uh(I,j,k)*dt   (j=js-2,je+2 ; I=Isq-2,Ieq+2)   and the same for vhtr
 * (src/core/MOM_dynamics_split_RK2.F90:1045-1052) on the resident UHTR / VHTR fields; uh, vh are Fortran-layout u- and
 * v-point arrays (host or device).  mom6adv_zero_transports replaces CS%uhtr(:,:,:) = 0.0 ; CS%vhtr(:,:,:) = 0.0
 * (src/core/MOM.F90:1465-1466).  mom6adv_column_diag replaces the vertical sums of post_tracer_transport_diagnostics
 * (src/tracer/MOM_tracer_registry.F90:725-815) for resident tracer m: kind ADV_XY_2D (:786-791, needs advection_xy
 * associated, h_diag ignored), ZINT (:796-802), ZINT_100M (:745-766,805-811); out2d is an (isd:ied, jsd:jed) array,
 * zero outside the computational domain like the reference's work2d. */
int mom6adv_accumulate_transports(mom6adv_plan *plan, const double *uh, const double *vh, double dt);
/* The same call with tracer_hordiff's diffusive-flux diagnostics (src/tracer/MOM_tracer_hor_diff.F90:388-405 zeroing,
 * :576-591 accumulation; Reg%Tr(m)%df_x, df_y, df2d_x, df2d_y of src/tracer/MOM_tracer_types.F90).
 * mom6adv_bind_tracer_df says which of the four are associated for tracer m (non-NULL = associated; the values live in HBM);
 * mom6adv_hordiff_diag runs the iteration and fills them (dt is tracer_hordiff's dt: Idt = 1/dt, :204);
 * mom6adv_download_diag fetches them with kind 5 df_x, 6 df_y, 7 df2d_x, 8 df2d_y.
 * STATE: built after the round's GPU time was spent -- the kernels have not run on a GPU.  The oracle for them
 * (ora_tracer_hordiff_df) is pinned against the translated reference lines; mom6adv_hordiff_diag returns
 * MOM6ADV_ERR_UNSUPPORTED unless the option "hordiff_df" is set. */
int mom6adv_bind_tracer_df(mom6adv_plan *plan, int m, const double *df_x, const double *df_y, const double *df2d_x,
                           const double *df2d_y);
int mom6adv_hordiff_diag(mom6adv_plan *plan, const double *h, const double *khdt_x, const double *khdt_y, int num_itts,
                         double h_neglect, int Diffuse_ML_interior, int nkml, int nk_rho_varies, double ML_KhTr_scale,
                         double dt);

/* mom6adv_offline_advect replaces the iteration loop of offline_advection_ale (SURVEY.md 8f rank 4;
 * src/tracer/MOM_offline_main.F90:254-312 set-up, :320-411 loop, :414-415 epilogue) with remaining_transport_sum
 * (:615-650) and its reproducing sum (src/framework/MOM_coms.F90:119-213) on the RESIDENT fields:
 *   in : H_END = h_pre, UHTR = uhtr, VHTR = vhtr (mom6adv_upload_field), the resident tracers
 *   out: H_END = h_pre (:414), UHR = uhtr and VHR = vhtr, the remaining transports after :415 (mom6adv_download_field),
 *        the tracers; *converged (:276, :400, :405), *n_iter = trips through the loop, tot_residual[0..*n_iter] =
 *        remaining_transport_sum at :312 and at :391 of every trip (NULL to skip)
 * Each trip calls advect_tracer with x_first_in = x_before_y (flipped every trip, :339), vol_prev = h_vol,
 * max_iter_in = 1, update_vol_prev = .true., uhr_out = uhtr, vhr_out = vhtr (:334-336).  dt_offline, num_off_iter,
 * off_ale_mod, x_before_y, min_residual are the offline_transport_CS members; H_to_kg_m2, kg_m2_to_H are GV's and L_to_m,
 * m_to_L are US's (:636, :647).  The ALE step inside the loop (:346-380) is other modules' work and is not built: the
 * call returns MOM6ADV_ERR_UNSUPPORTED if MODULO(iter, off_ale_mod) == 0 for any iter <= num_off_iter.  On a multi-block
 * plan the call is collective (group passes, an all-reduce of the sum's six integers).  Halos of the outputs are valid to
 * the plan's exchange depth. */
int mom6adv_offline_advect(mom6adv_plan *plan, double dt_offline, int num_off_iter, int off_ale_mod, int x_before_y,
                           double min_residual, double H_to_kg_m2, double L_to_m, double kg_m2_to_H, double m_to_L,
                           int *converged, int *n_iter, double *tot_residual);
/* mom6adv_hordiff replaces the iteration loop of tracer_hordiff, the call that follows advect_tracer in step_tracer
 * (src/core/MOM.F90:1441):
 *   do itt=1,num_itts ; do_group_pass(CS%pass_t) ; Coef_x/Coef_y/Ihdxdy ; dTr ; t = t + dTr ; conc_underflow ; enddo
 * (src/tracer/MOM_tracer_hor_diff.F90:539-607) on the plan's RESIDENT tracers (all bound tracers, as Reg%Tr(1:ntr)).
 *   h          layer thickness (isd:ied, jsd:jed, nk), host or device; NULL = the resident H_END field (step_tracer
 *              passes the same h to advect_tracer's h_end and to tracer_hordiff)
 *   khdt_x     (IsdB:IedB, jsd:jed) and khdt_y (isd:ied, JsdB:JedB): Khtr * dt * (face width / distance) as the
 *              reference's 2-D set-up leaves them (:186-366), host or device, valid on faces is-1..ie / js-1..je
 *   num_itts   :368-387 (the reference's max_across_PEs stays with the caller); I_numitts = 1/num_itts
 *   h_neglect  GV%H_subroundoff (:151)
 *   Diffuse_ML_interior, nkml, nk_rho_varies, ML_KhTr_scale   the level rule of :543-550
 * The df_x / df_y / df2d_x / df2d_y diagnostics (:576-591) are not produced: the call fails with
 * MOM6ADV_ERR_UNSUPPORTED rather than skip them silently if want_df is non-zero.  Epipycnal diffusion (:609-) and the
 * khdt set-up stay with the caller. */
int mom6adv_hordiff(mom6adv_plan *plan, const double *h, const double *khdt_x, const double *khdt_y, int num_itts,
                    double h_neglect, int Diffuse_ML_interior, int nkml, int nk_rho_varies, double ML_KhTr_scale,
                    int want_df);
int mom6adv_zero_transports(mom6adv_plan *plan);
enum { MOM6ADV_CD_ADV_XY_2D = 0, MOM6ADV_CD_ZINT = 1, MOM6ADV_CD_ZINT_100M = 2 };
int mom6adv_column_diag(mom6adv_plan *plan, int m, int kind, const double *h_diag, double H_to_m, double *out2d);

/* ---- no-allocation variant of the 100 m column integral ------------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arithmetic as
 * mom6adv_column_diag(kind = ZINT_100M) (src/tracer/MOM_tracer_registry.F90:745-766,805-811) without a plan:
 *   h_diag, t : DEVICE arrays (ni, nj, nk), Fortran order (i fastest), of the columns to integrate (the caller passes
 *               the computational domain); out2d : DEVICE array (ni, nj); stream : a cudaStream_t (NULL = default).
 * Calls no cudaMalloc / cudaFree / allocating routine: its scratch (per-CTA maxima and the layer bound khi) is a fixed
 * file-scope device array of 256 KiB + 256 B that every GPU loading the library holds until exit, which limits
 * ni*nj to MOM6ADV_NOALLOC_MAX_COLUMNS (beyond it: MOM6ADV_ERR_UNSUPPORTED and a message in errbuf).  Asynchronous on
 * `stream`.  All calls share the scratch: calls that overlap on the GPU race -- serialise them.  The first call on a
 * device resolves the scratch's device address (make it before starting a capture).  errbuf (nullable) receives the
 * error text. */
#define MOM6ADV_NOALLOC_MAX_COLUMNS 8388608LL
int mom6adv_zint100_noalloc(const double *h_diag, const double *t, int ni, int nj, int nk, double H_to_m,
                            double *out2d, void *stream, char *errbuf, int errbuf_len);

/* ---- observability ------------------------------------------------------- */
/* Device time [ms] of the last advect call, CUDA-event timed on the plan's
 * stream; the stand-in for the (Ocean advect tracer) clock (:1146). */
double mom6adv_last_advect_ms(const mom6adv_plan *plan);
/* Number of kernels the plan launched since creation. */
long long mom6adv_kernel_launches(const mom6adv_plan *plan);
/* Bytes of HBM held by the plan. */
long long mom6adv_device_bytes(const mom6adv_plan *plan);
/* Bytes the group passes (MOM_tracer_advect.F90:206) of this rank have listed for exchange with its neighbours since
 * plan creation (peer loads or NCCL messages; the dirty-tile skip may move fewer). */
long long mom6adv_halo_bytes(const mom6adv_plan *plan);
/* Options: "profile" 0/1 -- time every kernel class of advect calls with CUDA
 * events on the plan's stream (resets the accumulators); "fast" 0/1 -- 1 (default; any non-zero value) uses the
 * streaming kernels, flux diagnostics included, 0 forces the generic thread-per-cell sweeps (A/B timing);
 * "sparse" 0/1 -- 1 (default) updates only the cells next to faces with residual transport on trips >= 2;
 * "sparse_cap" n -- estimated list length above which a residual trip takes the dense kernels instead (0 = cells/16;
 * a hint: either choice gives the same bits);
 * "mirror_chunk" n -- tracers per pipeline stage of mom6adv_advect (default 2,
 * 0 = one stage: upload everything, compute, download everything);
 * "mirror_ahead" n -- uploads queued n chunks in front of the compute (default 2); "trace" 0/1
 * -- mom6adv_advect prints its copy-stream and compute times to stderr;
 * "pull" 0/1 -- multi-block plans: 1 (default) fills halos by loading the
 * neighbour GPU's fields directly (CUDA IPC + NVLink), 0 packs and uses NCCL
 * send/recv; choose before the first advect call; "pull2" 0/1 -- 1 (default): 16-byte pair-list peer loads;
 * "dirty" 0/1 -- 1 (default): after a list-driven trip the tracer pulls skip the tiles the owner has not changed;
 * "overlap" 0/1 -- 1 (default): on 1 x N layouts the tracer halo rows of the first pass travel under the x
 * sweep of the interior rows; "obc_multi" 0/1 -- 0 (default): mom6adv_set_obc refuses an associated OBC on a
 * multi-block plan; 1: every rank hands over its own clipped, locally indexed segments (segment%HI as
 * setup_segment_indices leaves it on that PE); each rank then equals the reference algorithm run per PE on the same
 * layout. */
int mom6adv_set_option(mom6adv_plan *plan, const char *key, int value);
/* Kernel classes of the per-class timers: the stand-ins for the reference's
 * (Ocean advect tracer) / (Ocean tracer halo updates) / (Ocean tracer global
 * synch) clock triple (:1146-1148), resolved per kernel family. */
enum { MOM6ADV_KC_PROLOGUE = 0, MOM6ADV_KC_HALO = 1, MOM6ADV_KC_FLAGS = 2, MOM6ADV_KC_MASS = 3,
       MOM6ADV_KC_TRACER_X = 4, MOM6ADV_KC_TRACER_Y = 5, MOM6ADV_KC_OTHER = 6, MOM6ADV_KC_SPARSE = 7,
       /* parts of HALO on a multi-block plan (already counted in HALO): */
       MOM6ADV_KC_EXCH_PACK = 8, MOM6ADV_KC_EXCH_WIRE = 9, MOM6ADV_KC_EXCH_UNPACK = 10,
       MOM6ADV_KC_COUNT = 11 };
/* Accumulated device time [ms] and number of timed brackets of one class
 * since "profile" was last set. */
int mom6adv_kernel_class_time(const mom6adv_plan *plan, int kclass, double *ms_total, long long *n);
/* The cudaStream_t every kernel of the plan is launched on (so that a host
 * can bracket calls with its own CUDA events). */
void *mom6adv_stream(const mom6adv_plan *plan);

#ifdef __cplusplus
}
#endif
#endif
