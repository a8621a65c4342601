// Host-callable launch wrappers of kernels.cu (internal).
#pragma once
#include "../../include/mom6adv.h"
#include "internal.cuh"

void launch_repack_in(cudaStream_t s, const Geom &g, double *dst, const double *src, int ilo, int jlo, int nlev);
void launch_repack_out(cudaStream_t s, const Geom &g, double *dst, const double *src, int ilo, int jlo, int nlev);
void launch_neglect(cudaStream_t s, const Geom &g, const double *areaT, const double *mCu, const double *mCv,
                    double Hsub, double *uhn, double *vhn, double *mmx, double *mmy);
void launch_init_residuals(cudaStream_t s, const Geom &g, const double *h_end, const double *uhtr,
                           const double *vhtr, const double *vol_prev, const double *areaT,
                           double *uhr, double *vhr, double *hprev);
void launch_rescan(cudaStream_t s, const Geom &g, const double *uhr, const double *vhr, const int *domore_k,
                   int *domore_u, int *domore_v, int isv, int iev, int jsv, int jev, int stencil);
void launch_domore_k(cudaStream_t s, const Geom &g, int *domore_k, const int *domore_u, const int *domore_v,
                     int ju0, int ju1, int jv0, int jv1);
void launch_sum_domore_k(cudaStream_t s, const int *domore_k, int nk, int *out);
void launch_flags_begin(cudaStream_t s, const Geom &g, const int *domore_k, int *domore, int *entry, int lo, int hi);
void launch_mass_flux(cudaStream_t s, int dir, const Geom &g, Range rg, const double *res, const double *hprev,
                      const double *areaT, double min_h, const int *entry, int *domore, double *hh, double *cfl,
                      double *hc, double *gq, int plm);
void launch_mass_update(cudaStream_t s, int dir, const Geom &g, Range rg, double *res, double *hprev,
                        const double *hh, const double *areaT, const double *neglect, double h_neglect,
                        const int *entry, const int *domore_k, double *hlst, double *ihnew,
                        int *list, int *count, int cap, int *nz = nullptr);
void launch_tracer_sweep(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TracerDev *tr, int ntr,
                         int src, const double *hh, const double *cfl, const double *hlst, const double *ihnew,
                         const double *mm, const double *IareaT, const int *entry, const int *domore_k, double Idt);
void launch_ad2d(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TracerDev *tr, int ntr, int src,
                 const double *hh, const double *cfl, const double *ihnew, const double *mm, const int *entry,
                 const int *domore_k, double Idt, const int *tflags = nullptr);
void launch_halo_apply(cudaStream_t s, long long plane, int nk, const int *dst, const int *src,
                       const signed char *sgn, int n, double *const *fields, int nfields);
void launch_halo_pack(cudaStream_t s, long long plane, int nk, const int *src, const signed char *sgn, int n,
                      double *const *fields, int nfields, double *buf);
void launch_halo_pull(cudaStream_t s, long long plane, long long peer_plane, int nk, const int *dst, const int *src,
                      const signed char *sgn, int n, double *const *fields, double *const *peer_fields, int nfields);
// the same exchange, 16 bytes per load and four levels in flight per thread (pair list: dst2/src2 address two
// consecutive points), skipping points whose source tile the owner has not changed since the last pass
// (dirty: a LOCAL copy of the owner's per-level tile map + its "everything changed" byte; nullptr: pull everything)
void launch_halo_pull2(cudaStream_t s, long long plane, long long peer_plane, int nk, const int *dst2, const int *src2,
                       const signed char *sgn2, const int *tile2, int n2, double *const *fields,
                       double *const *peer_fields, int nfields, const unsigned char *dirty, int peer_tiles_plane);
void launch_fetch_bytes(cudaStream_t s, unsigned char *dst, const unsigned char *src, long long bytes);
void launch_halo_unpack(cudaStream_t s, long long plane, int nk, const int *dst, int n,
                        double *const *fields, int nfields, const double *buf);
void launch_fill(cudaStream_t s, double *p, long long n, double v);
long long mass_tile_count(const Geom &g);
void launch_mass_flux_scan(cudaStream_t s, int dir, const Geom &g, Range rg, const double *res, const double *hprev,
                           const double *areaT, double min_h, const int *entry, int *domore, double *hh, double *cfl,
                           double *hc, double *gq, int plm, int *tflags, int *tlist, int *tcount, int nsm);
void launch_mass_update_list(cudaStream_t s, int dir, const Geom &g, Range rg, double *res, double *hprev,
                             const double *hh, const double *areaT, const double *neglect, double h_neglect,
                             const int *entry, const int *domore_k, double *hlst, double *ihnew,
                             int *list, int *count, int cap, const int *tlist, const int *tcount, int nsm);
void launch_fill_int(cudaStream_t s, int *p, int n, int v);
void launch_publish(cudaStream_t s, int *host_mapped, const int *dev, int n);

// fast.cu: streaming tracer sweeps
int x_fast_tiles(const Geom &g, int wbase);
void launch_x_fast(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *hc, const double *gq, const double *hlst, const double *ihnew, const double *mm, const int *entry,
                   const int *domore_k, int wbase, int wj_lo, int wj_hi,
                   const DiagTable *dg = nullptr, const double *IareaT = nullptr, double Idt = 0.0);
void launch_y_fast(cudaStream_t s, int scheme, const Geom &g, Range rg, const TrTable &tt, const double *hh,
                   const double *hc, const double *gq, const double *hlst, const double *ihnew, const double *mm, const int *entry,
                   const int *domore_k);
// ad_x | ad_y and advection_xy of the tracers that have them, from the sweep's INPUT values (kernels.cu): run before an
// in-place sweep.  Needs CFL itself (cfl), not the streaming kernels' factors.
void launch_flux_diag(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TracerDev *tr, int ntr, int src,
                      const double *hh, const double *cfl, const double *ihnew, const double *mm, const double *IareaT,
                      const int *entry, const int *domore_k, double Idt);
// sparse.cu: list-driven tracer update of the residual iterations
void launch_sparse_sweep(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TrTable &tt,
                         const int *list, const int *cnt_dev, int nsm, const double *hh, const double *cfl,
                         const double *hlst, const double *ihnew, const double *mm, const int *entry,
                         unsigned char *dirty, const DiagTable *dg = nullptr, const double *IareaT = nullptr,
                         double Idt = 0.0, const int *tflags = nullptr);

// step.cu: transport accumulation and column diagnostics either side of advect_tracer
void launch_accumulate(cudaStream_t s, const Geom &g, double *acc, const double *add, double dt, int ilo, int ihi,
                       int jlo, int jhi);
void launch_col_sum(cudaStream_t s, const Geom &g, const double *a, double *out);
void launch_col_zint(cudaStream_t s, const Geom &g, const double *h, const double *t, double H_to_m, double *out);
void launch_col_zint100(cudaStream_t s, const Geom &g, const double *h, const double *t, double H_to_m, int *khi,
                        double *out);

// halo.cu: exchange planning (host side)
#include <vector>
struct ExchangeSpec {
  int nig, njg, npi, npj, halo_i, halo_j;
  int cyclic_x, cyclic_y, tripolar_n, symmetric;
  int depth = 0;                                // halo points deeper than this are never read between two passes (0: all)
  std::vector<int> ibeg, iend, jbeg, jend;      // global compute extents of each layout column / row
};
struct HaloBuild { std::vector<int> dst, src; std::vector<signed char> sgn; };
struct PeerBuild {
  int rank = -1;
  std::vector<int> send_src[3], recv_dst[3];      // per stagger: 0 h, 1 u, 2 v
  std::vector<signed char> send_sgn[3];
  // peer-pull form of the same exchange: where each received point lives in the SENDER's padded layout
  std::vector<int> recv_src[3];
  std::vector<signed char> recv_sgn[3];
  long long peer_plane = 0;                       // the sender's plane size (its level stride)
  int peer_P = 0, peer_R = 0;                     // the sender's pitch and rows (dirty-tile map of its tracers)
};
void make_geom(const mom6adv_grid *gr, Geom &g, int &wbase);
// Is halo point (i,j) of stagger pos within `depth` ROWS of the computational domain (v faces: of a row that is)?
// Only whole rows are dropped: trimming columns would turn the 32-byte E/W halo pieces into partial-sector writes.
inline bool halo_point_used(int pos, int i, int j, int isc, int iec, int jsc, int jec, int depth) {
  (void)i; (void)isc; (void)iec;
  if (depth <= 0) return true;
  const int jlo = jsc - depth - (pos == 2);
  return j >= jlo && j <= jec + depth;
}
void build_halo_exchange(const Geom &g, const ExchangeSpec &x, int me, HaloBuild local[3], std::vector<PeerBuild> &peers);

// hordiff.cu: the diffusion iteration of tracer_hordiff (MOM_tracer_hor_diff.F90:539-607)
void launch_hd_coef(cudaStream_t s, const Geom &g, const HdParams &p, const double *h, const double *khdt_x,
                    const double *khdt_y, const double *IareaT, double *cx, double *cy, double *ihd);
void launch_hd_apply(cudaStream_t s, const Geom &g, const HdParams &p, const TrTable &tt, const double *cx,
                     const double *cy, const double *ihd);
void launch_hd_df(cudaStream_t s, const Geom &g, const HdParams &p, const TrTable &tt, const DfTable &df, bool any3,
                  bool any2, const double *cx, const double *cy, double Idt);

// offline.cu: the iteration loop of offline_advection_ale (MOM_offline_main.F90:254-415), element-wise parts
void launch_off_init(cudaStream_t s, const Geom &g, double *h);
void launch_off_halos(cudaStream_t s, const Geom &g, bool keep_w, bool keep_s, const double *hprev, const double *uhr,
                      const double *vhr, double *h, double *usub, double *vsub);
void launch_off_vol(cudaStream_t s, const Geom &g, const double *h, const double *areaT, double *vol);
void launch_off_post(cudaStream_t s, const Geom &g, const double *vol, const double *IareaT, double *h, const double *uhr,
                     const double *vhr, double *usub, double *vsub);
void launch_off_residual(cudaStream_t s, const Geom &g, const double *u, const double *v, const double *h,
                         const double *areaT, double Hsub, double HL2_to_kg_scale, unsigned long long *acc);
