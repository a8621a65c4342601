// C ABI of include/mom6adv.h: plan object, HBM residency, and the host-side
// iteration driver of advect_tracer (/root/reference/src/tracer/MOM_tracer_advect.F90:112-320).
#include "../../include/mom6adv.h"
#include "internal.cuh"
#include "kernels.h"
#include "obc.cuh"

#include <nccl.h>

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>
#include <limits>
#include <chrono>

#define MOM6ADV_VERSION "mom6adv-b200 0.1 (sm_100a)"

struct TracerSlot {
  double *host_t = nullptr;                 // MIRROR-mode location of Tr(m)%t (host or device, Fortran layout)
  double *host_diag[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};   // ad_x, ad_y, ad2d_x, ad2d_y, advection_xy
  double *buf[2] = {nullptr, nullptr};      // padded HBM storage (ping-pong)
  double *diag[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
  double *df[4] = {nullptr, nullptr, nullptr, nullptr};      // df_x, df_y, df2d_x, df2d_y of tracer_hordiff (mom6adv_bind_tracer_df)
  double conc_underflow = 0.0;
  bool bound = false;
  int cur = 0;                              // live ping-pong buffer
};

struct HaloList {       // one stagger position
  int n = 0;
  int *dst = nullptr, *src = nullptr;
  signed char *sgn = nullptr;
};

struct PeerLists {        // one neighbour rank of the multi-GPU group pass
  int rank = -1;
  int n_send[3] = {0, 0, 0}, n_recv[3] = {0, 0, 0};
  int *send_src[3] = {nullptr, nullptr, nullptr}, *recv_dst[3] = {nullptr, nullptr, nullptr};
  signed char *send_sgn[3] = {nullptr, nullptr, nullptr};
  double *sendbuf = nullptr, *recvbuf = nullptr;
  long long send_elems = 0, recv_elems = 0;
  // peer-pull exchange: source indices in the neighbour's layout and its fields mapped through CUDA IPC
  int *recv_src[3] = {nullptr, nullptr, nullptr};
  signed char *recv_sgn[3] = {nullptr, nullptr, nullptr};
  long long peer_plane = 0;
  // the same receive lists as 16-byte moves (two consecutive points per entry) + the leftover single points; tile = the
  // source's 32-column tile in the owner's plane (dirty-tile skip of the tracer pulls after a list-driven trip)
  int n_pair[3] = {0, 0, 0}, n_single[3] = {0, 0, 0};
  int *pair_dst[3] = {nullptr, nullptr, nullptr}, *pair_src[3] = {nullptr, nullptr, nullptr}, *pair_tile[3] = {nullptr, nullptr, nullptr};
  signed char *pair_sgn[3] = {nullptr, nullptr, nullptr};
  int *single_dst[3] = {nullptr, nullptr, nullptr}, *single_src[3] = {nullptr, nullptr, nullptr};
  signed char *single_sgn[3] = {nullptr, nullptr, nullptr};
  int peer_tiles_plane = 0;
  unsigned char *dirty_local = nullptr;      // this pass's copy of the owner's dirty map
  long long dirty_local_bytes = 0;
  std::vector<void *> mapped;        // cudaIpcOpenMemHandle results, slot order [uhr, vhr, hprev, buf0 x ntr, buf1 x ntr, dirty map]
  double **fields_dev = nullptr;     // two tables like halo_fields_dev, of the neighbour's pointers
};

#define MIRROR_CHUNK_DEFAULT 2
struct mom6adv_plan {
  Geom g;
  mom6adv_config cfg;
  mom6adv_domain dom;
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  double Angstrom_H = 0, H_subroundoff = 0;
  int IedB = 0, JedB = 0;
  // 2-D grid fields (one padded plane each)
  double *areaT = nullptr, *IareaT = nullptr, *mCu = nullptr, *mCv = nullptr;
  double *uh_neglect = nullptr, *vh_neglect = nullptr, *mmx = nullptr, *mmy = nullptr;
  // 3-D fields
  double *h_end = nullptr, *uhtr = nullptr, *vhtr = nullptr, *vol_prev = nullptr;
  double *uhr = nullptr, *vhr = nullptr, *hprev = nullptr;
  double *hh = nullptr, *cfl = nullptr, *hlst = nullptr, *ihnew = nullptr, *hc = nullptr, *gq = nullptr;
  int *domore_u = nullptr, *domore_v = nullptr, *entry_u = nullptr, *entry_v = nullptr, *domore_k = nullptr;
  int *sum_dev = nullptr, *sum_host = nullptr, *sum_host_dev = nullptr;   // sum_host: mapped pinned; sum_host_dev: its device alias
  double *staging = nullptr;
  unsigned long long *off_acc = nullptr;            // mom6adv_offline_advect: the six integers of its reproducing sum
  cudaEvent_t ev_off0 = nullptr, ev_off1 = nullptr;  // ... and the events around the whole loop (run_advect uses ev0 / ev1 inside it)
  double *khdt_x = nullptr, *khdt_y = nullptr;      // mom6adv_hordiff: 2-D diffusivity-times-dt fields (allocated on first use)
  double *stage_in = nullptr;        // MIRROR pipeline: ring of contiguous upload slots (slots x chunk x Fortran-layout tracer)
  long long stage_in_elems = 0;
  long long staging_elems = 0;
  std::vector<TracerSlot> tr;
  int ntr = 0;
  int cur = 0;                       // live ping-pong buffer of the tracers of the active range
  int m_lo = 0, m_hi = 0;            // active tracer range of the running advect call [m_lo, m_hi)
  cudaStream_t s_in = nullptr, s_out = nullptr;      // MIRROR mode: H2D / D2H copy streams
  std::vector<cudaEvent_t> ev_chunk;                 // per chunk c: [4c] uploaded, [4c+1] repacked, [4c+2] computed, [4c+3] downloaded
  cudaEvent_t ev_huv = nullptr;                      // h_end/uhtr/vhtr uploaded
  TracerDev *tr_dev = nullptr;       // device copy of the tracer table
  bool tr_dirty = true;
  double **halo_fields_dev = nullptr;  // [uhr], [vhr], [hprev, T...]
  int halo_fields_cap = 0;
  HaloList halo[3];                  // 0 = h points, 1 = u points, 2 = v points
  std::string err;
  double last_ms = 0.0;
  long long launches = 0;
  long long dev_bytes = 0;
  int opt_fused = 0;
  int opt_fast = 2;      // 0 generic thread-per-cell kernels (A/B timing, literal restatement), non-zero: streaming kernels
  // multi-GPU (LAYOUT = npi,npj): one rank per GPU, NCCL over NVLink
  bool multi = false;
  ncclComm_t comm = nullptr;
  int nranks = 1, rank = 0;
  std::vector<PeerLists> peers;
  int peer_ntr = -1;                 // ntr the exchange buffers were sized for
  long long halo_bytes_sent = 0;
  int opt_chunk = MIRROR_CHUNK_DEFAULT;  // tracers per pipeline stage of the MIRROR call
  int opt_ahead = 2;                 // chunks whose upload is queued in front of the running one
  int opt_pull = 1;                  // multi-GPU group pass by peer-memory loads (0: pack -> NCCL send/recv -> unpack)
  int ipc_ntr = -1;                  // registry size the IPC mappings were made for
  int *bar_dev = nullptr;            // scratch word of the neighbour barrier
  int *tflags = nullptr, *tlist = nullptr;   // per 32-face tile: outputs written / cells may change; tiles to update (residual trips)
  int nsm = 148;
  cudaStream_t s_pull = nullptr;     // tracer halos of the first pass are pulled here, under the interior x sweep
  cudaEvent_t ev_pull_go = nullptr, ev_pull_done = nullptr;
  bool pull_pending = false;         // the tracer halo rows are still in flight on s_pull
  int opt_overlap = 1;
  int opt_pull2 = 1;                 // 16-byte pair-list pull kernel (0: one point per thread)
  int opt_dirty = 1;                 // tracer pulls skip tiles the owner has not changed since the last pass
  int opt_sp_cap = 0;                // estimated list length above which a residual trip takes the dense kernels (0: cells/16)
  // open boundaries (mom6adv_set_obc)
  bool obc_active = false;
  bool obc_global = false;                  // some PE has an open boundary (the *_exist_globally flags): every rank keeps to dense trips
  int obc_flags[4] = {0, 0, 0, 0};          // open_u, open_v, specified_u, specified_v ..._BCs_exist_globally
  std::vector<ObcSeg> obc_segs;
  std::vector<ObcTr> obc_trs;
  std::vector<double *> obc_tres;           // device copies of the reservoirs
  ObcSeg *obc_segs_dev = nullptr;
  ObcTr *obc_trs_dev = nullptr;
  unsigned char *obc_cond = nullptr;
  long long obc_cond_bytes = 0;
  double *obc_ttmp = nullptr, *obc_slope = nullptr, *obc_fx = nullptr, *obc_hh2 = nullptr, *maskT = nullptr;
  int opt_trace = 0;                 // print the MIRROR pipeline's copy/compute times to stderr
  int opt_hordiff_df = 0;            // accept mom6adv_hordiff_diag (df_* diagnostics of tracer_hordiff: built, not yet run on a GPU)
  int opt_obc_multi = 0;             // experimental: accept open boundaries on a multi-block plan (see mom6adv_set_obc)
  int opt_sparse = 1;                // list-driven tracer update on residual iterations (sparse.cu)
  int *sp_list = nullptr; int sp_cap = 0;      // cells the residual trip updates (capacity: every cell, it cannot overflow)
  long long est_cells[2] = {-1, -1};      // host's estimate of the next trip's list length per direction (-1: unknown)
  bool est_dense[2] = {true, true};       // the count behind it is a dense trip's residual-face count (x2) / a list length
  unsigned char *dirty = nullptr;         // multi-GPU: per level and 32-column tile 'a tracer cell changed since the last pass' + one 'everything' byte
  long long dirty_bytes = 0;
  long long sparse_sweeps = 0, dense_sweeps = 0;
  int wbase = 0;                     // first column of the first x tile (fast.cu)
  // per-kernel-class CUDA-event timing (option "profile")
  int opt_profile = 0;
  std::vector<cudaEvent_t> prof_ev;      // pairs: [2n] start, [2n+1] stop
  std::vector<int> prof_cls;             // class of pair n (pairs used in the current call)
  double prof_ms[MOM6ADV_KC_COUNT] = {};
  long long prof_n[MOM6ADV_KC_COUNT] = {};
};

// RAII bracket: records an event pair around the launches of one kernel class.
struct ProfScope {
  mom6adv_plan *p; int idx = -1;
  ProfScope(mom6adv_plan *plan, int cls) : p(plan) {
    if (!p->opt_profile) return;
    idx = (int)p->prof_cls.size();
    while ((int)p->prof_ev.size() < 2 * (idx + 1)) { cudaEvent_t e; cudaEventCreate(&e); p->prof_ev.push_back(e); }
    p->prof_cls.push_back(cls);
    cudaEventRecord(p->prof_ev[2 * idx], p->stream);
  }
  ~ProfScope() { if (idx >= 0) cudaEventRecord(p->prof_ev[2 * idx + 1], p->stream); }
};
static void prof_collect(mom6adv_plan *p) {   // call after the stream is synchronised
  for (size_t n = 0; n < p->prof_cls.size(); n++) {
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, p->prof_ev[2 * n], p->prof_ev[2 * n + 1]) == cudaSuccess) {
      p->prof_ms[p->prof_cls[n]] += ms; p->prof_n[p->prof_cls[n]]++;
    }
  }
  p->prof_cls.clear();
}

static thread_local std::string g_err;
static void teardown_ipc(mom6adv_plan *plan, bool collective);
static void free_obc(mom6adv_plan *plan);

#define CK(call)                                                                            \
  do {                                                                                      \
    cudaError_t e_ = (call);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      char b_[512];                                                                         \
      snprintf(b_, sizeof b_, "CUDA error %s at %s:%d (%s)", cudaGetErrorString(e_), __FILE__, __LINE__, #call); \
      if (plan) plan->err = b_; g_err = b_;                                                 \
      return MOM6ADV_ERR_CUDA;                                                              \
    }                                                                                       \
  } while (0)

static int fail(mom6adv_plan *plan, int code, const std::string &msg) {
  if (plan) plan->err = msg;
  g_err = msg;
  return code;
}

static long long field_elems(const mom6adv_plan *p) { return p->g.plane * p->g.nk; }

template <typename T>
static int dev_alloc(mom6adv_plan *plan, T **out, long long n) {
  CK(cudaMalloc((void **)out, sizeof(T) * (size_t)std::max<long long>(n, 1)));
  CK(cudaMemsetAsync(*out, 0, sizeof(T) * (size_t)std::max<long long>(n, 1), plan->stream));
  plan->dev_bytes += (long long)sizeof(T) * n;
  return 0;
}

static bool is_device_ptr(const void *p) {
  cudaPointerAttributes a;
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

// Fortran-layout array (ilo:ied, jlo:jed, nlev) -> padded field
static int upload(mom6adv_plan *plan, double *dst, const double *src, int ilo, int jlo, int nlev) {
  const Geom &g = plan->g;
  const long long n = (long long)(g.ied - ilo + 1) * (g.jed - jlo + 1) * nlev;
  if (is_device_ptr(src)) {
    launch_repack_in(plan->stream, g, dst, src, ilo, jlo, nlev);
  } else {
    CK(cudaMemcpyAsync(plan->staging, src, sizeof(double) * n, cudaMemcpyHostToDevice, plan->stream));
    launch_repack_in(plan->stream, g, dst, plan->staging, ilo, jlo, nlev);
  }
  plan->launches++;
  CK(cudaGetLastError());
  return 0;
}
static int download(mom6adv_plan *plan, double *dst, const double *src, int ilo, int jlo, int nlev) {
  const Geom &g = plan->g;
  const long long n = (long long)(g.ied - ilo + 1) * (g.jed - jlo + 1) * nlev;
  if (is_device_ptr(dst)) {
    launch_repack_out(plan->stream, g, dst, src, ilo, jlo, nlev);
  } else {
    launch_repack_out(plan->stream, g, plan->staging, src, ilo, jlo, nlev);
    CK(cudaMemcpyAsync(dst, plan->staging, sizeof(double) * n, cudaMemcpyDeviceToHost, plan->stream));
  }
  plan->launches++;
  CK(cudaGetLastError());
  return 0;
}

// ---------------------------------------------------------------------------
// Halo lists: for every halo point of a stagger, the computational-domain
// point that a full-width group pass would copy into it (closed boundaries
// have none).  Same index algebra as FMS's update for cyclic edges and the
// north fold (scalars mirror in i; C-grid vector components also change sign).
// Single-block version (npi = npj = 1); the multi-GPU lists are built by
// halo.cu from the same locate() rule.
// ---------------------------------------------------------------------------
struct HostHalo { std::vector<int> dst, src; std::vector<signed char> sgn; };

static bool locate_self(const mom6adv_plan *p, int pos, int ig, int jg, int *il, int *jl, int *sg) {
  const mom6adv_domain &d = p->dom;
  const Geom &g = p->g;
  *sg = 1;
  if (d.tripolar_n && jg > d.njg) {
    if (pos == 0) { ig = d.nig + 1 - ig; jg = 2 * d.njg + 1 - jg; }
    else if (pos == 1) { ig = d.nig - ig; jg = 2 * d.njg + 1 - jg; *sg = -1; }
    else { ig = d.nig + 1 - ig; jg = 2 * d.njg - jg; *sg = -1; }
  }
  const int ilo = (pos == 1) ? 0 : 1, jlo = (pos == 2) ? 0 : 1;
  if (d.cyclic_x) { while (ig < 1) ig += d.nig; while (ig > d.nig) ig -= d.nig; }
  else if (ig < ilo || ig > d.nig) return false;
  if (d.cyclic_y) { while (jg < 1) jg += d.njg; while (jg > d.njg) jg -= d.njg; }
  else if (jg < jlo || jg > d.njg) return false;
  int i = ig - d.ioff, j = jg - d.joff;
  int i0 = g.isc, j0 = g.jsc;
  if (pos == 1 && ig == 0) i0 = g.isc - 1;
  if (pos == 2 && jg == 0) j0 = g.jsc - 1;
  if (i >= i0 && i <= g.iec && j >= j0 && j <= g.jec) { *il = i; *jl = j; return true; }
  return false;
}

// Halo depth that can be read between two group passes: nsten_halo*stencil (:208-210).  Deeper halo points
// (the outermost row when halo = 4 and stencil = 3) are not exchanged: the reference fills them, nothing reads them.
static int exch_depth(const mom6adv_plan *p) {
  const Geom &g = p->g;
  const int stencil = (p->cfg.usePPM && !p->cfg.useHuynhStencilBug) ? 3 : 2;
  const int halo_min = std::min(std::min(g.isc - g.isd, g.ied - g.iec), std::min(g.jsc - g.jsd, g.jed - g.jec));
  return std::max(1, halo_min / stencil) * stencil;
}

static void build_halo_host(const mom6adv_plan *p, int pos, HostHalo &out) {
  const Geom &g = p->g;
  const bool symmetric = (g.IsdB == g.isd - 1);
  const int ilo = (pos == 1) ? g.IsdB : g.isd, jlo = (pos == 2) ? g.JsdB : g.jsd;
  int ic0 = g.isc, jc0 = g.jsc;
  if (symmetric) { if (pos == 1) ic0 = g.isc - 1; if (pos == 2) jc0 = g.jsc - 1; }
  for (int j = jlo; j <= g.jed; j++) for (int i = ilo; i <= g.ied; i++) {
    if (i >= ic0 && i <= g.iec && j >= jc0 && j <= g.jec) continue;
    if (!halo_point_used(pos, i, j, g.isc, g.iec, g.jsc, g.jec, exch_depth(p))) continue;
    int il, jl, sg;
    if (!locate_self(p, pos, i + p->dom.ioff, j + p->dom.joff, &il, &jl, &sg)) continue;
    out.dst.push_back((i + g.coff) + g.P * (j + g.roff));
    out.src.push_back((il + g.coff) + g.P * (jl + g.roff));
    out.sgn.push_back((signed char)sg);
  }
}

static int upload_halo_list(mom6adv_plan *plan, const HostHalo &h, HaloList &d) {
  d.n = (int)h.dst.size();
  if (d.n == 0) return 0;
  CK(cudaMalloc((void **)&d.dst, sizeof(int) * d.n));
  CK(cudaMalloc((void **)&d.src, sizeof(int) * d.n));
  CK(cudaMalloc((void **)&d.sgn, d.n));
  CK(cudaMemcpy(d.dst, h.dst.data(), sizeof(int) * d.n, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.src, h.src.data(), sizeof(int) * d.n, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(d.sgn, h.sgn.data(), d.n, cudaMemcpyHostToDevice));
  plan->dev_bytes += 9LL * d.n;
  return 0;
}

static int create_impl(mom6adv_plan *plan, const mom6adv_grid *gr) {
  Geom &g = plan->g;
  CK(cudaSetDevice(plan->device));
  CK(cudaStreamCreateWithFlags(&plan->stream, cudaStreamNonBlocking));
  CK(cudaEventCreate(&plan->ev0));
  CK(cudaEventCreate(&plan->ev1));
  const long long pl = g.plane;
  double **p2[] = {&plan->areaT, &plan->IareaT, &plan->mCu, &plan->mCv,
                   &plan->uh_neglect, &plan->vh_neglect, &plan->mmx, &plan->mmy, &plan->maskT};
  for (double **pp : p2) { int rc = dev_alloc(plan, pp, pl); if (rc) return rc; }
  const long long fe = field_elems(plan);
  double **p3[] = {&plan->h_end, &plan->uhtr, &plan->vhtr, &plan->uhr, &plan->vhr, &plan->hprev,
                   &plan->hh, &plan->cfl, &plan->hlst, &plan->ihnew, &plan->hc, &plan->gq};
  for (double **pp : p3) { int rc = dev_alloc(plan, pp, fe); if (rc) return rc; }
  int **pf[] = {&plan->domore_u, &plan->domore_v, &plan->entry_u, &plan->entry_v};
  for (int **pp : pf) { int rc = dev_alloc(plan, pp, (long long)g.R * g.nk); if (rc) return rc; }
  { int rc = dev_alloc(plan, &plan->domore_k, g.nk); if (rc) return rc; }
  { int rc = dev_alloc(plan, &plan->sum_dev, 8); if (rc) return rc; }
  CK(cudaHostAlloc((void **)&plan->sum_host, sizeof(int) * 8, cudaHostAllocMapped));
  CK(cudaHostGetDevicePointer((void **)&plan->sum_host_dev, plan->sum_host, 0));
  // staging buffer: the largest Fortran-layout 3-D field (u or v stagger)
  plan->staging_elems = (long long)(g.ied - g.isd + 2) * (g.jed - g.jsd + 2) * g.nk;
  { int rc = dev_alloc(plan, &plan->staging, plan->staging_elems); if (rc) return rc; }

  // grid fields
  int rc;
  if ((rc = upload(plan, plan->areaT, gr->areaT, g.isd, g.jsd, 1))) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  if ((rc = upload(plan, plan->IareaT, gr->IareaT, g.isd, g.jsd, 1))) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  if (gr->mask2dT) { if ((rc = upload(plan, plan->maskT, gr->mask2dT, g.isd, g.jsd, 1))) return rc; CK(cudaStreamSynchronize(plan->stream)); }
  if ((rc = upload(plan, plan->mCu, gr->mask2dCu, g.IsdB, g.jsd, 1))) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  if ((rc = upload(plan, plan->mCv, gr->mask2dCv, g.isd, g.JsdB, 1))) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  launch_neglect(plan->stream, g, plan->areaT, plan->mCu, plan->mCv, plan->H_subroundoff,
                 plan->uh_neglect, plan->vh_neglect, plan->mmx, plan->mmy);
  plan->launches++;
  CK(cudaGetLastError());

  if (!plan->multi) {
    for (int pos = 0; pos < 3; pos++) {
      HostHalo hh;
      build_halo_host(plan, pos, hh);
      if ((rc = upload_halo_list(plan, hh, plan->halo[pos]))) return rc;
    }
  }
  CK(cudaStreamSynchronize(plan->stream));
  return 0;
}

#define CKN(call)                                                                           \
  do {                                                                                      \
    ncclResult_t e_ = (call);                                                               \
    if (e_ != ncclSuccess) {                                                                \
      char b_[512];                                                                         \
      snprintf(b_, sizeof b_, "NCCL error %s at %s:%d (%s)", ncclGetErrorString(e_), __FILE__, __LINE__, #call); \
      if (plan) plan->err = b_; g_err = b_;                                                 \
      return MOM6ADV_ERR_COMM;                                                              \
    }                                                                                       \
  } while (0)

template <typename T>
static int upload_vec(mom6adv_plan *plan, const std::vector<T> &h, T **d) {
  *d = nullptr;
  if (h.empty()) return 0;
  CK(cudaMalloc((void **)d, sizeof(T) * h.size()));
  CK(cudaMemcpy(*d, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice));
  plan->dev_bytes += (long long)(sizeof(T) * h.size());
  return 0;
}

extern "C" int mom6adv_comm_unique_id(char *id128) {
  if (!id128) return MOM6ADV_ERR_INVALID;
  ncclUniqueId id;
  if (ncclGetUniqueId(&id) != ncclSuccess) return fail(nullptr, MOM6ADV_ERR_COMM, "ncclGetUniqueId failed");
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
  memcpy(id128, &id, 128);
  return MOM6ADV_OK;
}

// Exchange lists of this rank (host side of comm_init; also what mom6adv_exchange_list reports).
static void exchange_spec(const Geom &g, const mom6adv_domain &d, const int *ibeg, const int *iend, const int *jbeg,
                          const int *jend, ExchangeSpec &x) {
  x.nig = d.nig; x.njg = d.njg; x.npi = d.npi; x.npj = d.npj;
  x.halo_i = g.isc - g.isd; x.halo_j = g.jsc - g.jsd;
  x.cyclic_x = d.cyclic_x; x.cyclic_y = d.cyclic_y; x.tripolar_n = d.tripolar_n;
  x.symmetric = (g.IsdB == g.isd - 1);
  x.ibeg.assign(ibeg, ibeg + d.npi); x.iend.assign(iend, iend + d.npi);
  x.jbeg.assign(jbeg, jbeg + d.npj); x.jend.assign(jend, jend + d.npj);
}

extern "C" int mom6adv_comm_init(mom6adv_plan *plan, const char *id128, const int *ibeg, const int *iend,
                                 const int *jbeg, const int *jend) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "comm_init: plan not initialised");
  if (!ibeg || !iend || !jbeg || !jend) return fail(plan, MOM6ADV_ERR_INVALID, "comm_init: NULL extents");
  if (plan->comm || !plan->peers.empty() || plan->halo[0].n || plan->halo[1].n || plan->halo[2].n)
    return fail(plan, MOM6ADV_ERR_INVALID, "comm_init: called twice");
  CK(cudaSetDevice(plan->device));
  {
    // the exchange lists are built for every rank from the layout alone (halo.cu rank_geo): local indexing with the data
    // domain starting at 1 (config_src/infra/FMS2/MOM_domain_infra.F90:1815-1821) and equal halos on both sides
    const Geom &g0 = plan->g;
    if (g0.isd != 1 || g0.jsd != 1 || (g0.isc - g0.isd) != (g0.ied - g0.iec) || (g0.jsc - g0.jsd) != (g0.jed - g0.jec))
      return fail(plan, MOM6ADV_ERR_UNSUPPORTED, "comm_init: multi-block plans need local indexing (isd = jsd = 1) and symmetric halos");
  }
  ExchangeSpec x;
  exchange_spec(plan->g, plan->dom, ibeg, iend, jbeg, jend, x);
  x.depth = exch_depth(plan);
  if (x.ibeg[plan->dom.pi] != plan->g.isc + plan->dom.ioff || x.jbeg[plan->dom.pj] != plan->g.jsc + plan->dom.joff ||
      x.iend[plan->dom.pi] != plan->g.iec + plan->dom.ioff || x.jend[plan->dom.pj] != plan->g.jec + plan->dom.joff)
    return fail(plan, MOM6ADV_ERR_INVALID, "comm_init: extents disagree with this rank's grid");
  HaloBuild local[3];
  std::vector<PeerBuild> pb;
  build_halo_exchange(plan->g, x, plan->rank, local, pb);
  int rc;
  for (int pos = 0; pos < 3; pos++) {
    HostHalo hh; hh.dst = local[pos].dst; hh.src = local[pos].src; hh.sgn = local[pos].sgn;
    if ((rc = upload_halo_list(plan, hh, plan->halo[pos]))) return rc;
  }
  for (const PeerBuild &b : pb) {
    PeerLists pl;
    pl.rank = b.rank;
    for (int pos = 0; pos < 3; pos++) {
      pl.n_send[pos] = (int)b.send_src[pos].size(); pl.n_recv[pos] = (int)b.recv_dst[pos].size();
      if ((rc = upload_vec(plan, b.send_src[pos], &pl.send_src[pos]))) return rc;
      if ((rc = upload_vec(plan, b.send_sgn[pos], &pl.send_sgn[pos]))) return rc;
      if ((rc = upload_vec(plan, b.recv_dst[pos], &pl.recv_dst[pos]))) return rc;
      if ((rc = upload_vec(plan, b.recv_src[pos], &pl.recv_src[pos]))) return rc;
      if ((rc = upload_vec(plan, b.recv_sgn[pos], &pl.recv_sgn[pos]))) return rc;
    }
    pl.peer_plane = b.peer_plane;
    pl.peer_tiles_plane = (b.peer_P >> 5) * b.peer_R;
    for (int pos = 0; pos < 3; pos++) {
      std::vector<int> pd, ps, pt, sd, ss; std::vector<signed char> pg, sg;
      const std::vector<int> &D = b.recv_dst[pos], &S = b.recv_src[pos];
      const std::vector<signed char> &G = b.recv_sgn[pos];
      for (size_t t = 0; t < D.size();) {
        const bool pair = t + 1 < D.size() && D[t + 1] == D[t] + 1 && S[t + 1] == S[t] + 1 && G[t + 1] == G[t] &&
                          (D[t] & 1) == 0 && (S[t] & 1) == 0 && b.peer_P > 0;
        if (pair) {
          const int rj = S[t] / b.peer_P, ci = S[t] - rj * b.peer_P;
          pd.push_back(D[t]); ps.push_back(S[t]); pg.push_back(G[t]); pt.push_back(rj * (b.peer_P >> 5) + (ci >> 5));
          t += 2;
        } else { sd.push_back(D[t]); ss.push_back(S[t]); sg.push_back(G[t]); t += 1; }
      }
      pl.n_pair[pos] = (int)pd.size(); pl.n_single[pos] = (int)sd.size();
      if ((rc = upload_vec(plan, pd, &pl.pair_dst[pos]))) return rc;
      if ((rc = upload_vec(plan, ps, &pl.pair_src[pos]))) return rc;
      if ((rc = upload_vec(plan, pg, &pl.pair_sgn[pos]))) return rc;
      if ((rc = upload_vec(plan, pt, &pl.pair_tile[pos]))) return rc;
      if ((rc = upload_vec(plan, sd, &pl.single_dst[pos]))) return rc;
      if ((rc = upload_vec(plan, ss, &pl.single_src[pos]))) return rc;
      if ((rc = upload_vec(plan, sg, &pl.single_sgn[pos]))) return rc;
    }
    plan->peers.push_back(pl);
  }
  if (id128 && plan->nranks > 1) {
    ncclUniqueId id;
    memcpy(&id, id128, 128);
    CKN(ncclCommInitRank(&plan->comm, plan->nranks, id, plan->rank));
  }
  return MOM6ADV_OK;
}

// Host-side view of the exchange plan (tests; no device work).  what: 0 local dst, 1 local src, 2 local sign,
// 3 send src to `peer`, 4 send sign to `peer`, 5 recv dst from `peer`, 6 / 7 the same points' index in `peer`'s own
// padded layout and their sign (peer-pull form).  Returns the count (<0: error);
// fills at most cap entries.  geom (nullable) receives P, R, coff, roff of the padded layout.
extern "C" int mom6adv_exchange_list(const mom6adv_grid *gr, const mom6adv_domain *dom, const int *ibeg,
                                     const int *iend, const int *jbeg, const int *jend, int peer, int pos, int what,
                                     int *out, int cap, int *geom) {
  if (!gr || !dom || pos < 0 || pos > 2) return -1;
  Geom g; int wbase;
  make_geom(gr, g, wbase);
  ExchangeSpec x;
  exchange_spec(g, *dom, ibeg, iend, jbeg, jend, x);
  HaloBuild local[3];
  std::vector<PeerBuild> pb;
  build_halo_exchange(g, x, dom->pi + dom->npi * dom->pj, local, pb);
  if (geom) { geom[0] = g.P; geom[1] = g.R; geom[2] = g.coff; geom[3] = g.roff; }
  std::vector<int> v;
  if (what == 0) v = local[pos].dst;
  else if (what == 1) v = local[pos].src;
  else if (what == 2) v.assign(local[pos].sgn.begin(), local[pos].sgn.end());
  else {
    for (const PeerBuild &b : pb) if (b.rank == peer) {
      if (what == 3) v = b.send_src[pos];
      else if (what == 4) v.assign(b.send_sgn[pos].begin(), b.send_sgn[pos].end());
      else if (what == 5) v = b.recv_dst[pos];
      else if (what == 6) v = b.recv_src[pos];
      else if (what == 7) v.assign(b.recv_sgn[pos].begin(), b.recv_sgn[pos].end());
    }
  }
  for (int n = 0; n < (int)v.size() && n < cap; n++) out[n] = v[n];
  return (int)v.size();
}

// Padded HBM layout of one plane (internal.cuh) from the Fortran index bounds.
void make_geom(const mom6adv_grid *gr, Geom &g, int &wbase) {
  const int hi = gr->isc - gr->isd;
  g.isd = gr->isd; g.ied = gr->ied; g.jsd = gr->jsd; g.jed = gr->jed;
  g.isc = gr->isc; g.iec = gr->iec; g.jsc = gr->jsc; g.jec = gr->jec;
  g.IsdB = gr->IsdB; g.JsdB = gr->JsdB; g.nk = gr->nk;
  const int nid = g.ied - g.isd + 1, njd = g.jed - g.jsd + 1;
  // column of isd; isc lands on a multiple of 32 = the y sweep's column tile, so that a sweep over isc..iec does not
  // spend a half-used tile at either end (OM4_025: 45 tiles instead of 46)
  int lead = ((hi + 1 + 31) / 32) * 32 - hi;
  if (lead < 4) lead += 32;
  g.coff = lead - g.isd;
  // x tiles of the streaming kernel: 128 columns from wbase + 120*n, 120 useful (lanes 1..30)
  wbase = (lead - 4) & ~3;
  const int ntx = (lead + nid - 1 - (wbase + 4)) / 120 + 1;
  g.P = ((std::max(lead + nid + 1, wbase + 120 * (ntx - 1) + 128) + 63) / 64) * 64;   // y tiles are 64 columns
  g.roff = 1 - g.jsd;
  g.R = njd + 2;
  g.plane = (long long)g.P * g.R;
}

extern "C" const char *mom6adv_version(void) { return MOM6ADV_VERSION; }

extern "C" const char *mom6adv_last_error(const mom6adv_plan *plan) {
  return plan ? plan->err.c_str() : g_err.c_str();
}

extern "C" int mom6adv_plan_create(const mom6adv_grid *gr, const mom6adv_config *c, const mom6adv_domain *dom,
                                   int device, mom6adv_plan **out) {
  if (!gr || !c || !out) return fail(nullptr, MOM6ADV_ERR_INVALID, "plan_create: NULL argument");
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return fail(nullptr, MOM6ADV_ERR_CUDA, "plan_create: no CUDA device (this library has no CPU fallback)");
  }
  if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) return fail(nullptr, MOM6ADV_ERR_CUDA, "cudaGetDevice failed"); }
  if (device >= ndev) return fail(nullptr, MOM6ADV_ERR_INVALID, "plan_create: bad device index");
  if (!gr->areaT || !gr->IareaT || !gr->mask2dCu || !gr->mask2dCv)
    return fail(nullptr, MOM6ADV_ERR_INVALID, "plan_create: grid arrays missing");
  const int hi = gr->isc - gr->isd, hj = gr->jsc - gr->jsd;
  if (gr->nk < 1 || gr->iec < gr->isc || gr->jec < gr->jsc || hi < 0 || hj < 0 ||
      gr->IsdB < gr->isd - 1 || gr->IsdB > gr->isd || gr->JsdB < gr->jsd - 1 || gr->JsdB > gr->jsd)
    return fail(nullptr, MOM6ADV_ERR_INVALID, "plan_create: inconsistent grid bounds");

  mom6adv_plan *plan = new mom6adv_plan();
  plan->cfg = *c;
  plan->device = device;
  plan->Angstrom_H = gr->Angstrom_H;
  plan->H_subroundoff = gr->H_subroundoff;
  plan->IedB = gr->IedB; plan->JedB = gr->JedB;
  Geom &g = plan->g;
  make_geom(gr, g, plan->wbase);
  if (dom) plan->dom = *dom;
  else {
    mom6adv_domain d = {};
    d.nig = g.iec - g.isc + 1; d.njg = g.jec - g.jsc + 1; d.npi = d.npj = 1;
    d.ioff = -(g.isc - 1); d.joff = -(g.jsc - 1);
    plan->dom = d;
  }
  plan->multi = (plan->dom.npi * plan->dom.npj != 1);     // lists and communicator come from mom6adv_comm_init
  plan->nranks = plan->dom.npi * plan->dom.npj;
  plan->rank = plan->dom.pi + plan->dom.npi * plan->dom.pj;
  int rc = create_impl(plan, gr);
  if (rc) { g_err = plan->err; mom6adv_plan_destroy(plan); return rc; }
  *out = plan;
  return MOM6ADV_OK;
}

static void free_slot(TracerSlot &s) {
  for (int b = 0; b < 2; b++) if (s.buf[b]) { cudaFree(s.buf[b]); s.buf[b] = nullptr; }
  for (int d = 0; d < 5; d++) if (s.diag[d]) { cudaFree(s.diag[d]); s.diag[d] = nullptr; }
  for (int d = 0; d < 4; d++) if (s.df[d]) { cudaFree(s.df[d]); s.df[d] = nullptr; }
}

extern "C" int mom6adv_plan_destroy(mom6adv_plan *plan) {
  if (!plan) return MOM6ADV_OK;
  cudaSetDevice(plan->device);
  if (plan->stream) cudaStreamSynchronize(plan->stream);
  teardown_ipc(plan, true);
  free_obc(plan);
  { double *o[] = {plan->obc_ttmp, plan->obc_slope, plan->obc_fx, plan->obc_hh2, plan->maskT}; for (double *p : o) if (p) cudaFree(p); }
  if (plan->bar_dev) cudaFree(plan->bar_dev);
  if (plan->tflags) cudaFree(plan->tflags);
  if (plan->tlist) cudaFree(plan->tlist);
  double *d[] = {plan->areaT, plan->IareaT, plan->mCu, plan->mCv, plan->uh_neglect, plan->vh_neglect, plan->mmx,
                 plan->mmy, plan->h_end, plan->uhtr, plan->vhtr, plan->vol_prev, plan->uhr, plan->vhr, plan->hprev,
                 plan->hh, plan->cfl, plan->hlst, plan->ihnew, plan->hc, plan->gq, plan->staging};
  for (double *p : d) if (p) cudaFree(p);
  int *f[] = {plan->domore_u, plan->domore_v, plan->entry_u, plan->entry_v, plan->domore_k, plan->sum_dev};
  for (int *p : f) if (p) cudaFree(p);
  if (plan->sum_host) cudaFreeHost(plan->sum_host);
  for (PeerLists &p : plan->peers) {
    for (int pos = 0; pos < 3; pos++) { if (p.send_src[pos]) cudaFree(p.send_src[pos]); if (p.send_sgn[pos]) cudaFree(p.send_sgn[pos]); if (p.recv_dst[pos]) cudaFree(p.recv_dst[pos]); if (p.recv_src[pos]) cudaFree(p.recv_src[pos]); if (p.recv_sgn[pos]) cudaFree(p.recv_sgn[pos]); }
    for (int pos = 0; pos < 3; pos++) {
      void *q[] = {p.pair_dst[pos], p.pair_src[pos], p.pair_tile[pos], p.pair_sgn[pos], p.single_dst[pos], p.single_src[pos], p.single_sgn[pos]};
      for (void *x : q) if (x) cudaFree(x);
    }
    if (p.dirty_local) cudaFree(p.dirty_local);
    if (p.sendbuf) cudaFree(p.sendbuf);
    if (p.recvbuf) cudaFree(p.recvbuf);
  }
  if (plan->comm) ncclCommDestroy(plan->comm);
  if (plan->sp_list) cudaFree(plan->sp_list);
  if (plan->off_acc) cudaFree(plan->off_acc);
  if (plan->ev_off0) { cudaEventDestroy(plan->ev_off0); cudaEventDestroy(plan->ev_off1); }
  if (plan->khdt_x) cudaFree(plan->khdt_x);
  if (plan->khdt_y) cudaFree(plan->khdt_y);
  if (plan->dirty) cudaFree(plan->dirty);
  for (auto &s : plan->tr) free_slot(s);
  if (plan->tr_dev) cudaFree(plan->tr_dev);
  if (plan->halo_fields_dev) cudaFree(plan->halo_fields_dev);
  for (int p = 0; p < 3; p++) {
    if (plan->halo[p].dst) cudaFree(plan->halo[p].dst);
    if (plan->halo[p].src) cudaFree(plan->halo[p].src);
    if (plan->halo[p].sgn) cudaFree(plan->halo[p].sgn);
  }
  for (cudaEvent_t e : plan->prof_ev) cudaEventDestroy(e);
  for (cudaEvent_t e : plan->ev_chunk) cudaEventDestroy(e);
  if (plan->stage_in) cudaFree(plan->stage_in);
  if (plan->ev_huv) cudaEventDestroy(plan->ev_huv);
  if (plan->s_pull) { cudaStreamDestroy(plan->s_pull); cudaEventDestroy(plan->ev_pull_go); cudaEventDestroy(plan->ev_pull_done); }
  if (plan->s_in) cudaStreamDestroy(plan->s_in);
  if (plan->s_out) cudaStreamDestroy(plan->s_out);
  if (plan->ev0) cudaEventDestroy(plan->ev0);
  if (plan->ev1) cudaEventDestroy(plan->ev1);
  if (plan->stream) cudaStreamDestroy(plan->stream);
  delete plan;
  return MOM6ADV_OK;
}

// ---------------------------------------------------------------------------
// OBC argument of advect_tracer (:64).  Copies the segment tables and reservoirs to the device.
// ---------------------------------------------------------------------------
static void free_obc(mom6adv_plan *plan) {
  for (double *p : plan->obc_tres) if (p) cudaFree(p);
  plan->obc_tres.clear(); plan->obc_segs.clear(); plan->obc_trs.clear();
  if (plan->obc_segs_dev) { cudaFree(plan->obc_segs_dev); plan->obc_segs_dev = nullptr; }
  if (plan->obc_trs_dev) { cudaFree(plan->obc_trs_dev); plan->obc_trs_dev = nullptr; }
  if (plan->obc_cond) { cudaFree(plan->obc_cond); plan->obc_cond = nullptr; }
  plan->obc_cond_bytes = 0;
  plan->obc_active = false;
}

extern "C" int mom6adv_set_obc(mom6adv_plan *plan, const mom6adv_obc *obc) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "set_obc: plan not initialised");
  CK(cudaSetDevice(plan->device));
  CK(cudaStreamSynchronize(plan->stream));
  free_obc(plan);
  plan->obc_global = obc && (obc->open_u_BCs_exist_globally || obc->open_v_BCs_exist_globally ||
                             obc->specified_u_BCs_exist_globally || obc->specified_v_BCs_exist_globally);
  if (!obc || !obc->OBC_pe) return MOM6ADV_OK;                 // unassociated / no segment on this PE (:442, :580)
  // Multi-block plans: refused unless the option "obc_multi" is set.  With it every rank hands over ITS segments,
  // clipped to its data domain and in its local indices, exactly as setup_segment_indices leaves segment%HI on each PE
  // (MOM_open_boundary.F90:1296-1328); a rank whose segments are all off its tile passes OBC_pe = false and takes the
  // plain sweeps.  Each rank then equals the reference algorithm run per PE on the same layout.
  if (plan->multi && !plan->opt_obc_multi)
    return fail(plan, MOM6ADV_ERR_UNSUPPORTED, "set_obc: open boundaries on a multi-block plan need the option obc_multi (every rank then passes its own clipped, locally indexed segments)");
  if (obc->number_of_segments < 0 || (obc->number_of_segments > 0 && !obc->segment))
    return fail(plan, MOM6ADV_ERR_INVALID, "set_obc: bad segment list");
  const Geom &g = plan->g;
  plan->obc_flags[0] = obc->open_u_BCs_exist_globally; plan->obc_flags[1] = obc->open_v_BCs_exist_globally;
  plan->obc_flags[2] = obc->specified_u_BCs_exist_globally; plan->obc_flags[3] = obc->specified_v_BCs_exist_globally;
  long long cond_bytes = 0;
  for (int n = 0; n < obc->number_of_segments; n++) {
    const mom6adv_obc_segment &h = obc->segment[n];
    ObcSeg sg;
    sg.dir = h.direction; sg.is_ew = h.is_E_or_W; sg.is_ns = h.is_N_or_S; sg.specified = h.specified;
    sg.has_reg = (h.ntseg >= 0);
    sg.isd = h.isd; sg.ied = h.ied; sg.jsd = h.jsd; sg.jed = h.jed;
    sg.IsdB = h.IsdB; sg.IedB = h.IedB; sg.JsdB = h.JsdB; sg.JedB = h.JedB;
    sg.ntseg = std::max(h.ntseg, 0); sg.tr_off = (int)plan->obc_trs.size();
    if (!sg.is_ew && !sg.is_ns) {      // a segment that is not on this PE (MOM_open_boundary.F90:1443-1447 return before the flags are set): keeps its number, does nothing
      sg.has_reg = false; sg.ntseg = 0; sg.cond_off = cond_bytes;
      plan->obc_segs.push_back(sg);
      continue;
    }
    if (sg.is_ew == sg.is_ns) return fail(plan, MOM6ADV_ERR_INVALID, "set_obc: a segment must be either E/W or N/S");
    // the faces and the cells next to them must lie inside the data domain (the reference indexes them unchecked)
    const bool ok = sg.is_ew ? (sg.IsdB - 2 >= g.isd && sg.IsdB + 3 <= g.ied && sg.jsd >= g.jsd && sg.jed <= g.jed && sg.jsd <= sg.jed)
                             : (sg.JsdB - 2 >= g.jsd && sg.JsdB + 3 <= g.jed && sg.isd >= g.isd && sg.ied <= g.ied && sg.isd <= sg.ied && sg.JedB >= sg.JsdB);
    if (!ok) return fail(plan, MOM6ADV_ERR_INVALID, "set_obc: segment outside the data domain");
    sg.cond_off = cond_bytes;
    const long long nfaces = sg.is_ew ? (long long)(sg.jed - sg.jsd + 1) : (long long)(sg.ied - sg.isd + 1) * (sg.JedB - sg.JsdB + 1);
    const long long ntres = sg.is_ew ? (long long)(sg.IedB - sg.IsdB + 1) * (sg.jed - sg.jsd + 1) * g.nk : nfaces * g.nk;
    cond_bytes += nfaces * g.nk;
    for (int e = 0; e < sg.ntseg; e++) {
      const mom6adv_obc_tracer &t = h.Tr[e];
      if (t.ntr_index < 1 || t.ntr_index > plan->ntr) return fail(plan, MOM6ADV_ERR_INVALID, "set_obc: ntr_index outside the registry (call set_ntr first)");
      ObcTr d; d.m = t.ntr_index - 1; d.conc = t.OBC_inflow_conc; d.tres = nullptr;
      if (t.tres) {
        double *dev = nullptr;
        CK(cudaMalloc((void **)&dev, sizeof(double) * ntres));
        plan->obc_tres.push_back(dev);
        if (is_device_ptr(t.tres)) CK(cudaMemcpyAsync(dev, t.tres, sizeof(double) * ntres, cudaMemcpyDeviceToDevice, plan->stream));
        else CK(cudaMemcpyAsync(dev, t.tres, sizeof(double) * ntres, cudaMemcpyHostToDevice, plan->stream));
        d.tres = dev;
      }
      plan->obc_trs.push_back(d);
    }
    plan->obc_segs.push_back(sg);
  }
  CK(cudaMalloc((void **)&plan->obc_segs_dev, sizeof(ObcSeg) * std::max<size_t>(plan->obc_segs.size(), 1)));
  CK(cudaMalloc((void **)&plan->obc_trs_dev, sizeof(ObcTr) * std::max<size_t>(plan->obc_trs.size(), 1)));
  if (!plan->obc_segs.empty()) CK(cudaMemcpyAsync(plan->obc_segs_dev, plan->obc_segs.data(), sizeof(ObcSeg) * plan->obc_segs.size(), cudaMemcpyHostToDevice, plan->stream));
  if (!plan->obc_trs.empty()) CK(cudaMemcpyAsync(plan->obc_trs_dev, plan->obc_trs.data(), sizeof(ObcTr) * plan->obc_trs.size(), cudaMemcpyHostToDevice, plan->stream));
  plan->obc_cond_bytes = std::max<long long>(cond_bytes, 1);
  CK(cudaMalloc((void **)&plan->obc_cond, plan->obc_cond_bytes));
  const long long fe = field_elems(plan);
  double **scr[] = {&plan->obc_ttmp, &plan->obc_slope, &plan->obc_fx, &plan->obc_hh2};
  for (double **pp : scr) if (!*pp) { int rc = dev_alloc(plan, pp, fe); if (rc) return rc; }
  CK(cudaStreamSynchronize(plan->stream));
  plan->obc_active = true;
  return MOM6ADV_OK;
}

// One sweep with an associated OBC (obc.cu): the reference's order of operations, tracer by tracer.
static int sweep_obc(mom6adv_plan *plan, int dir, Range rg, double Idt, int scheme) {
  const Geom &g = plan->g;
  cudaStream_t s = plan->stream;
  const double min_h = 0.1 * plan->Angstrom_H, h_neglect = plan->H_subroundoff;
  const bool x = (dir == 0);
  int *dom = x ? plan->domore_u : plan->domore_v, *ent = x ? plan->entry_u : plan->entry_v;
  double *res = x ? plan->uhr : plan->vhr;
  const double *mm = x ? plan->mmx : plan->mmy;
  const long long fe = field_elems(plan);
  const int open_g = plan->obc_flags[x ? 0 : 1], spec_g = plan->obc_flags[x ? 2 : 3];
  {
    ProfScope ps(plan, MOM6ADV_KC_MASS);
    launch_flags_begin(s, g, plan->domore_k, dom, ent, x ? rg.js : rg.js - 1, rg.je);
    launch_mass_flux(s, dir, g, rg, res, plan->hprev, plan->areaT, min_h, ent, dom, plan->hh, plan->cfl, nullptr, nullptr, 0);
    CK(cudaMemcpyAsync(plan->obc_hh2, plan->hh, sizeof(double) * fe, cudaMemcpyDeviceToDevice, s));
    CK(cudaMemsetAsync(plan->obc_cond, 0, plan->obc_cond_bytes, s));
    // :580-628 / :965-1014 -- uhh := uhr on the faces that take a prescribed inflow
    if (spec_g || open_g)
      for (const ObcSeg &sg : plan->obc_segs) {
        if (!sg.has_reg || (x ? !sg.is_ew : !sg.is_ns)) continue;
        if (!x && !sg.specified) continue;                    // the y form tests segment%specified here (:970), the x form does not
        launch_obc_mass(s, dir, g, rg, sg, 1, res, plan->obc_hh2, plan->maskT, plan->obc_cond, ent, plan->domore_k);
        plan->launches++;
      }
    if (open_g)
      for (const ObcSeg &sg : plan->obc_segs) {
        if (!sg.has_reg || sg.specified || (x ? !sg.is_ew : !sg.is_ns)) continue;
        launch_obc_mass(s, dir, g, rg, sg, 2, res, plan->obc_hh2, plan->maskT, plan->obc_cond, ent, plan->domore_k);
        plan->launches++;
      }
    launch_mass_update(s, dir, g, rg, res, plan->hprev, plan->obc_hh2, plan->areaT, x ? plan->uh_neglect : plan->vh_neglect,
                       h_neglect, ent, plan->domore_k, plan->hlst, plan->ihnew, nullptr, nullptr, 0);
    plan->launches += 3;
  }
  ProfScope pt(plan, x ? MOM6ADV_KC_TRACER_X : MOM6ADV_KC_TRACER_Y);
  const int sten = (scheme == SCHEME_PPM_CW) ? 2 : 1;           // the sweeps' local stencil (:398-399 / :770-771)
  for (int m = plan->m_lo; m < plan->m_hi; m++) {
    const TracerSlot &ts = plan->tr[m];
    TracerDev td;
    td.buf[0] = ts.buf[0]; td.buf[1] = ts.buf[1];
    td.ad_x = ts.diag[0]; td.ad_y = ts.diag[1]; td.ad2d_x = ts.diag[2]; td.ad2d_y = ts.diag[3]; td.adv_xy = ts.diag[4];
    td.conc_underflow = ts.conc_underflow;
    const double *T = ts.buf[plan->cur];
    CK(cudaMemcpyAsync(plan->obc_ttmp, T, sizeof(double) * fe, cudaMemcpyDeviceToDevice, s));     // T_tmp (:436-440 / :820-822)
    if (scheme != SCHEME_PPM_H3) launch_obc_slope(s, dir, g, rg, sten, T, mm, plan->obc_slope, plan->domore_k);
    for (const ObcSeg &sg : plan->obc_segs) {                   // :442-478 / :825-862, in segment order
      if (!sg.has_reg || (x ? !sg.is_ew : !sg.is_ns)) continue;
      launch_obc_override(s, dir, g, rg, sg, plan->obc_trs_dev, m, plan->obc_ttmp, plan->obc_slope, mm, ent, plan->domore_k);
      plan->launches++;
    }
    double *a2 = ts.diag[x ? 2 : 3];
    launch_obc_flux_update(s, scheme, dir, g, rg, td, m, plan->cur, plan->obc_ttmp, plan->obc_slope, plan->hh, plan->obc_hh2,
                           plan->cfl, plan->hlst, plan->ihnew, mm, plan->IareaT, ent, plan->domore_k, Idt,
                           plan->obc_segs_dev, (int)plan->obc_segs.size(), plan->obc_trs_dev, plan->obc_cond, a2 ? plan->obc_fx : nullptr);
    if (a2) { launch_obc_ad2d(s, dir, g, rg, a2, plan->obc_fx, plan->ihnew, ent, plan->domore_k, Idt); plan->launches++; }
    plan->launches += 2;
  }
  // x is out of place; y too, but its result goes back into the buffer it read, so that the live buffer alternates
  // exactly as on the streaming path (x flips, y in place): on a multi-block plan the ranks without a segment take
  // that path, and the neighbours' pulls address the live buffer by the common count of flips
  if (x) plan->cur = 1 - plan->cur;
  else
    for (int m = plan->m_lo; m < plan->m_hi; m++)
      CK(cudaMemcpyAsync(plan->tr[m].buf[plan->cur], plan->tr[m].buf[1 - plan->cur], sizeof(double) * fe, cudaMemcpyDeviceToDevice, s));
  CK(cudaGetLastError());
  return 0;
}

extern "C" int mom6adv_set_ntr(mom6adv_plan *plan, int ntr) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "set_ntr: plan not initialised");
  if (ntr < 0) return fail(plan, MOM6ADV_ERR_INVALID, "set_ntr: ntr < 0");
  CK(cudaSetDevice(plan->device));
  for (int m = ntr; m < (int)plan->tr.size(); m++) free_slot(plan->tr[m]);
  plan->tr.resize(ntr);
  const long long fe = field_elems(plan);
  for (int m = 0; m < ntr; m++)
    for (int b = 0; b < 2; b++)
      if (!plan->tr[m].buf[b]) { int rc = dev_alloc(plan, &plan->tr[m].buf[b], fe); if (rc) return rc; }
  plan->ntr = ntr;
  plan->tr_dirty = true;
  return MOM6ADV_OK;
}

extern "C" int mom6adv_bind_tracer(mom6adv_plan *plan, int m, double *t, double conc_underflow, double *ad_x,
                                   double *ad_y, double *ad2d_x, double *ad2d_y, double *advection_xy) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "bind_tracer: plan not initialised");
  if (m < 0 || m >= plan->ntr) return fail(plan, MOM6ADV_ERR_INVALID, "bind_tracer: tracer index out of range");
  CK(cudaSetDevice(plan->device));
  TracerSlot &s = plan->tr[m];
  s.host_t = t;
  s.conc_underflow = conc_underflow;
  double *hd[5] = {ad_x, ad_y, ad2d_x, ad2d_y, advection_xy};
  for (int d = 0; d < 5; d++) {
    s.host_diag[d] = hd[d];
    const bool two_d = (d == 2 || d == 3);
    if (hd[d] && !s.diag[d]) { int rc = dev_alloc(plan, &s.diag[d], two_d ? plan->g.plane : field_elems(plan)); if (rc) return rc; }
    if (!hd[d] && s.diag[d]) { cudaFree(s.diag[d]); s.diag[d] = nullptr; }
  }
  s.bound = true;
  plan->tr_dirty = true;
  return MOM6ADV_OK;
}

// The diffusive-flux diagnostics of tracer_hordiff (MOM_tracer_types.F90: df_x, df_y, df2d_x, df2d_y) of tracer m: a
// non-NULL argument says "associated" -- the values live in HBM (mom6adv_hordiff_diag fills them, mom6adv_download_diag
// kinds 5..8 fetch them); NULL frees the field.
extern "C" int mom6adv_bind_tracer_df(mom6adv_plan *plan, int m, const double *df_x, const double *df_y,
                                      const double *df2d_x, const double *df2d_y) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "bind_tracer_df: plan not initialised");
  if (m < 0 || m >= plan->ntr) return fail(plan, MOM6ADV_ERR_INVALID, "bind_tracer_df: tracer index out of range");
  CK(cudaSetDevice(plan->device));
  TracerSlot &s = plan->tr[m];
  const double *hd[4] = {df_x, df_y, df2d_x, df2d_y};
  for (int d = 0; d < 4; d++) {
    if (hd[d] && !s.df[d]) { int rc = dev_alloc(plan, &s.df[d], d >= 2 ? plan->g.plane : field_elems(plan)); if (rc) return rc; }
    if (!hd[d] && s.df[d]) { cudaFree(s.df[d]); s.df[d] = nullptr; }
  }
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

// Enable a diagnostic in RESIDENT mode without a host array (kind as in download_diag).
static int sync_tracer_table(mom6adv_plan *plan) {
  if (!plan->tr_dirty) return 0;
  std::vector<TracerDev> h(std::max(plan->ntr, 1));
  for (int m = 0; m < plan->ntr; m++) {
    const TracerSlot &s = plan->tr[m];
    h[m].buf[0] = s.buf[0]; h[m].buf[1] = s.buf[1];
    h[m].ad_x = s.diag[0]; h[m].ad_y = s.diag[1]; h[m].ad2d_x = s.diag[2]; h[m].ad2d_y = s.diag[3]; h[m].adv_xy = s.diag[4];
    h[m].conc_underflow = s.conc_underflow;
  }
  if (plan->ipc_ntr >= 0) plan->ipc_ntr = -2;          // buffers may have moved: map again
  if (plan->tr_dev) { cudaFree(plan->tr_dev); plan->tr_dev = nullptr; }
  CK(cudaMalloc((void **)&plan->tr_dev, sizeof(TracerDev) * h.size()));
  CK(cudaMemcpyAsync(plan->tr_dev, h.data(), sizeof(TracerDev) * h.size(), cudaMemcpyHostToDevice, plan->stream));
  // halo field tables, one per ping-pong buffer b: [0] = uhr, [1] = vhr, [2] = hprev, [3..] = tracers (buffer b)
  const int nf = 3 + plan->ntr;
  std::vector<double *> hf(2 * nf);
  for (int b = 0; b < 2; b++) {
    hf[b * nf + 0] = plan->uhr; hf[b * nf + 1] = plan->vhr; hf[b * nf + 2] = plan->hprev;
    for (int m = 0; m < plan->ntr; m++) hf[b * nf + 3 + m] = plan->tr[m].buf[b];
  }
  if (plan->halo_fields_dev) { cudaFree(plan->halo_fields_dev); plan->halo_fields_dev = nullptr; }
  CK(cudaMalloc((void **)&plan->halo_fields_dev, sizeof(double *) * 2 * nf));
  CK(cudaMemcpyAsync(plan->halo_fields_dev, hf.data(), sizeof(double *) * 2 * nf, cudaMemcpyHostToDevice, plan->stream));
  CK(cudaStreamSynchronize(plan->stream));
  plan->tr_dirty = false;
  return 0;
}

// ---------------------------------------------------------------------------
// do_group_pass (:206) on one device.
// ---------------------------------------------------------------------------
// ---------------------------------------------------------------------------
// Peer-pull exchange.  Every rank exports its exchanged buffers (uhr, vhr, hprev, both buffers of every
// tracer) as CUDA IPC handles, the handles travel in one ncclAllGather, and each rank maps its neighbours'
// buffers.  A group pass is then a gather kernel per neighbour that loads the owner's values over NVLink
// and stores them into the local halo -- no pack buffer, no send/recv, no unpack -- fenced by two
// one-word all-reduces (data ready / done reading).  Falls back to the NCCL send/recv path on all ranks
// together if any mapping fails.
// ---------------------------------------------------------------------------
static int exch_barrier(mom6adv_plan *plan) {
  CKN(ncclAllReduce(plan->bar_dev, plan->bar_dev, 1, ncclInt, ncclMax, plan->comm, plan->stream));
  return 0;
}

static void teardown_ipc(mom6adv_plan *plan, bool collective) {
  bool any = false;
  for (PeerLists &p : plan->peers) {
    for (void *q : p.mapped) if (q) { cudaIpcCloseMemHandle(q); any = true; }
    p.mapped.clear();
    if (p.fields_dev) { cudaFree(p.fields_dev); p.fields_dev = nullptr; }
  }
  // nobody may free a buffer a neighbour still has mapped
  if (any && collective && plan->comm && plan->bar_dev) {
    ncclAllReduce(plan->bar_dev, plan->bar_dev, 1, ncclInt, ncclMax, plan->comm, plan->stream);
    cudaStreamSynchronize(plan->stream);
  }
  plan->ipc_ntr = -1;
}

static int setup_ipc(mom6adv_plan *plan) {
  if (!plan->multi || plan->peers.empty() || !plan->comm || !plan->opt_pull) return 0;
  if (plan->ipc_ntr == plan->ntr) return 0;
  cudaStream_t s = plan->stream;
  if (!plan->bar_dev) { int rc = dev_alloc(plan, &plan->bar_dev, 4); if (rc) return rc; }
  teardown_ipc(plan, true);
  const int ntr = plan->ntr, nslot = 3 + 2 * ntr + 1, nr = plan->nranks;
  if (!plan->dirty) {      // tile map + the "everything" byte, padded so that readers fetch whole 16-byte words
    plan->dirty_bytes = (((plan->g.plane >> 5) * plan->g.nk + 16 + 15) / 16) * 16 + 16;
    int rc = dev_alloc(plan, &plan->dirty, plan->dirty_bytes); if (rc) return rc;
  }
  std::vector<cudaIpcMemHandle_t> mine(nslot), all((size_t)nslot * nr);
  std::vector<void *> slot(nslot);
  slot[0] = plan->uhr; slot[1] = plan->vhr; slot[2] = plan->hprev;
  for (int m = 0; m < ntr; m++) { slot[3 + m] = plan->tr[m].buf[0]; slot[3 + ntr + m] = plan->tr[m].buf[1]; }
  slot[3 + 2 * ntr] = plan->dirty;
  int ok = 1;
  for (int q = 0; q < nslot; q++)
    if (!slot[q] || cudaIpcGetMemHandle(&mine[q], slot[q]) != cudaSuccess) { cudaGetLastError(); ok = 0; memset(&mine[q], 0, sizeof mine[q]); }
  unsigned char *d_mine = nullptr, *d_all = nullptr;
  const size_t hb = sizeof(cudaIpcMemHandle_t) * nslot;
  CK(cudaMalloc((void **)&d_mine, hb)); CK(cudaMalloc((void **)&d_all, hb * nr));
  CK(cudaMemcpyAsync(d_mine, mine.data(), hb, cudaMemcpyHostToDevice, s));
  CKN(ncclAllGather(d_mine, d_all, hb, ncclChar, plan->comm, s));
  CK(cudaMemcpyAsync(all.data(), d_all, hb * nr, cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  cudaFree(d_mine); cudaFree(d_all);
  for (PeerLists &p : plan->peers) {
    p.mapped.assign(nslot, nullptr);
    for (int q = 0; q < nslot && ok; q++)
      if (cudaIpcOpenMemHandle(&p.mapped[q], all[(size_t)p.rank * nslot + q], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
        cudaGetLastError(); p.mapped[q] = nullptr; ok = 0;
      }
  }
  // all ranks take the same path
  CK(cudaMemcpyAsync(plan->bar_dev + 1, &ok, sizeof(int), cudaMemcpyHostToDevice, s));
  CKN(ncclAllReduce(plan->bar_dev + 1, plan->bar_dev + 1, 1, ncclInt, ncclMin, plan->comm, s));
  CK(cudaMemcpyAsync(&ok, plan->bar_dev + 1, sizeof(int), cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  if (!ok) {
    teardown_ipc(plan, true);
    plan->opt_pull = 0;
    if (plan->rank == 0) fprintf(stderr, "[mom6adv] CUDA IPC mapping of neighbour fields failed on some rank: group pass uses NCCL send/recv\n");
    return 0;
  }
  const int nf = 3 + ntr;
  for (PeerLists &p : plan->peers) {
    std::vector<double *> hf(2 * nf);
    for (int b = 0; b < 2; b++) {
      hf[b * nf + 0] = (double *)p.mapped[0]; hf[b * nf + 1] = (double *)p.mapped[1]; hf[b * nf + 2] = (double *)p.mapped[2];
      for (int m = 0; m < ntr; m++) hf[b * nf + 3 + m] = (double *)p.mapped[3 + b * ntr + m];
    }
    CK(cudaMalloc((void **)&p.fields_dev, sizeof(double *) * 2 * nf));
    CK(cudaMemcpy(p.fields_dev, hf.data(), sizeof(double *) * 2 * nf, cudaMemcpyHostToDevice));
  }
  plan->ipc_ntr = ntr;
  return 0;
}

enum { PASS_ALL = 0, PASS_TRACERS = 1, PASS_MASS = 2 };
static int group_pass(mom6adv_plan *plan, bool first_of_call, bool defer_tracers, int what = PASS_ALL) {
  const Geom &g = plan->g;
  cudaStream_t s = plan->stream;
  ProfScope ps(plan, MOM6ADV_KC_HALO);
  // u list -> uhr, v list -> vhr, h list -> hprev + tracers (live buffer)
  // PASS_MASS: uhr, vhr, hprev without the tracers (the offline loop's pass_var / pass_vector, MOM_offline_main.F90:386-387)
  const int nact = (what == PASS_MASS) ? 0 : plan->m_hi - plan->m_lo;
  double **tab = plan->halo_fields_dev + plan->cur * (3 + plan->ntr);
  // field-table segments by stagger: h = hprev + the active tracers, u = uhr, v = vhr
  struct Seg { double **t; int n; };
  // tracers_only: CS%pass_t of tracer_hordiff (MOM_tracer_hor_diff.F90:540) -- no hprev / uhr / vhr
  const int nm = (what == PASS_TRACERS) ? 0 : 1;
  const Seg segs[3][2] = {{{tab + 2, nm}, {tab + 3 + plan->m_lo, nact}}, {{tab + 0, nm}, {nullptr, 0}}, {{tab + 1, nm}, {nullptr, 0}}};
  // (1) halo points whose source lies in this rank's own block (cyclic wrap / fold onto itself)
  for (int pos = 0; pos < 3; pos++) for (const Seg &sg : segs[pos]) if (sg.n > 0) {
    launch_halo_apply(s, g.plane, g.nk, plan->halo[pos].dst, plan->halo[pos].src, plan->halo[pos].sgn, plan->halo[pos].n, sg.t, sg.n);
    plan->launches++;
  }
  if (plan->multi && !plan->peers.empty()) {
    if (!plan->comm) return fail(plan, MOM6ADV_ERR_COMM, "advect: mom6adv_comm_init has not been called on this multi-block plan");
    if (plan->opt_pull && plan->ipc_ntr == plan->ntr) {
      // the neighbours' previous writes must be complete: at the first pass of a call a barrier says so, later
      // passes follow the all-reduce of the convergence flags, which every rank enters after its sweeps
      if (first_of_call) { int rc = exch_barrier(plan); if (rc) return rc; }
      // q = 0: hprev / uhr / vhr, q = 1: the tracers.  With defer_tracers the tracer rows travel on s_pull while the
      // x sweep of the interior rows runs (1 x N layouts: those rows need no neighbour data); the sweep of the halo
      // rows waits for ev_pull_done (sweep()).  Tracer sources are the neighbours' live buffers, which nobody
      // writes before the convergence all-reduce of this trip, so only the mass fields need the closing fence.
      const bool use_dirty = !first_of_call && !defer_tracers && plan->opt_dirty && plan->opt_pull2 && plan->dirty;
      auto pull = [&](cudaStream_t st, int q) {
        for (PeerLists &p : plan->peers) {
          double **ptab = p.fields_dev + plan->cur * (3 + plan->ntr);
          const Seg psegs[3][2] = {{{ptab + 2, nm}, {ptab + 3 + plan->m_lo, nact}}, {{ptab + 0, nm}, {nullptr, 0}}, {{ptab + 1, nm}, {nullptr, 0}}};
          // the tracers' pull skips tiles the owner has not touched since the last pass (after a list-driven trip:
          // nearly all of the strip); the first pass of a call and the mass fields (3 of 55) are pulled whole
          const unsigned char *pd = nullptr;
          if (q == 1 && use_dirty) {
            if (!p.dirty_local) {
              p.dirty_local_bytes = (((long long)p.peer_tiles_plane * g.nk + 16 + 15) / 16) * 16;
              if (cudaMalloc((void **)&p.dirty_local, (size_t)p.dirty_local_bytes) != cudaSuccess) { cudaGetLastError(); p.dirty_local = nullptr; }
            }
            if (p.dirty_local) {
              launch_fetch_bytes(st, p.dirty_local, (const unsigned char *)p.mapped[3 + 2 * plan->ntr], p.dirty_local_bytes);
              pd = p.dirty_local;
              plan->launches++;
            }
          }
          for (int pos = 0; pos < 3; pos++) if (segs[pos][q].n > 0 && !plan->opt_pull2) {
            launch_halo_pull(st, g.plane, p.peer_plane, g.nk, p.recv_dst[pos], p.recv_src[pos], p.recv_sgn[pos], p.n_recv[pos],
                             segs[pos][q].t, psegs[pos][q].t, segs[pos][q].n);
            plan->launches++;
          }
          for (int pos = 0; pos < 3; pos++) if (segs[pos][q].n > 0 && plan->opt_pull2) {
            launch_halo_pull2(st, g.plane, p.peer_plane, g.nk, p.pair_dst[pos], p.pair_src[pos], p.pair_sgn[pos], p.pair_tile[pos],
                              p.n_pair[pos], segs[pos][q].t, psegs[pos][q].t, segs[pos][q].n, pd, p.peer_tiles_plane);
            launch_halo_pull(st, g.plane, p.peer_plane, g.nk, p.single_dst[pos], p.single_src[pos], p.single_sgn[pos], p.n_single[pos],
                             segs[pos][q].t, psegs[pos][q].t, segs[pos][q].n);
            plan->launches += 2;
          }
        }
      };
      for (PeerLists &p : plan->peers)
        plan->halo_bytes_sent += 8LL * g.nk * ((long long)(nm + nact) * p.n_recv[0] + nm * (p.n_recv[1] + p.n_recv[2]));
      { ProfScope pw(plan, MOM6ADV_KC_EXCH_WIRE);
        pull(s, 0);
        if (!defer_tracers) pull(s, 1);
      }
      if (defer_tracers) {
        if (!plan->s_pull) {
          int prio_lo = 0, prio_hi = 0;
          CK(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
          CK(cudaStreamCreateWithPriority(&plan->s_pull, cudaStreamNonBlocking, prio_hi));   // its blocks go ahead of the sweep's
          CK(cudaEventCreateWithFlags(&plan->ev_pull_go, cudaEventDisableTiming));
          CK(cudaEventCreateWithFlags(&plan->ev_pull_done, cudaEventDisableTiming));
        }
        CK(cudaEventRecord(plan->ev_pull_go, s));
        CK(cudaStreamWaitEvent(plan->s_pull, plan->ev_pull_go, 0));
        pull(plan->s_pull, 1);
        CK(cudaEventRecord(plan->ev_pull_done, plan->s_pull));
        plan->pull_pending = true;
      }
      CK(cudaGetLastError());
      // nobody overwrites its rows before every neighbour has read them
      { int rc = exch_barrier(plan); if (rc) return rc; }
      // every neighbour has read this rank's dirty map: start the next interval clean
      if (plan->dirty) CK(cudaMemsetAsync(plan->dirty, 0, (size_t)plan->dirty_bytes, s));
      return 0;
    }
    // (2) pack one message per neighbour: [h: hprev, tracers][u: uhr][v: vhr], each [field][level][point]
    { ProfScope pp(plan, MOM6ADV_KC_EXCH_PACK);
    for (PeerLists &p : plan->peers) {
      long long off = 0;
      for (int pos = 0; pos < 3; pos++) for (const Seg &sg : segs[pos]) if (sg.n > 0) {
        launch_halo_pack(s, g.plane, g.nk, p.send_src[pos], p.send_sgn[pos], p.n_send[pos], sg.t, sg.n, p.sendbuf + off);
        off += (long long)sg.n * g.nk * p.n_send[pos];
        plan->launches++;
      }
    }
    }
    CK(cudaGetLastError());
    // (3) device-resident exchange over NVLink
    { ProfScope pw(plan, MOM6ADV_KC_EXCH_WIRE);
    CKN(ncclGroupStart());
    for (PeerLists &p : plan->peers) {
      const long long ns = (long long)g.nk * ((long long)(nm + nact) * p.n_send[0] + nm * (p.n_send[1] + p.n_send[2]));
      const long long nr = (long long)g.nk * ((long long)(nm + nact) * p.n_recv[0] + nm * (p.n_recv[1] + p.n_recv[2]));
      if (ns) CKN(ncclSend(p.sendbuf, (size_t)ns, ncclDouble, p.rank, plan->comm, s));
      if (nr) CKN(ncclRecv(p.recvbuf, (size_t)nr, ncclDouble, p.rank, plan->comm, s));
      plan->halo_bytes_sent += 8 * ns;
    }
    CKN(ncclGroupEnd());
    }
    // (4) unpack
    ProfScope pu(plan, MOM6ADV_KC_EXCH_UNPACK);
    for (PeerLists &p : plan->peers) {
      long long off = 0;
      for (int pos = 0; pos < 3; pos++) for (const Seg &sg : segs[pos]) if (sg.n > 0) {
        launch_halo_unpack(s, g.plane, g.nk, p.recv_dst[pos], p.n_recv[pos], sg.t, sg.n, p.recvbuf + off);
        off += (long long)sg.n * g.nk * p.n_recv[pos];
        plan->launches++;
      }
    }
  }
  CK(cudaGetLastError());
  return 0;
}

// Exchange buffers are sized by ntr; (re)allocate when the registry changed.
static int size_peer_buffers(mom6adv_plan *plan) {
  if (!plan->multi || plan->peer_ntr == plan->ntr) return 0;
  if (plan->opt_pull && plan->ipc_ntr == plan->ntr) return 0;      // the pull path needs no message buffers
  const Geom &g = plan->g;
  for (PeerLists &p : plan->peers) {
    if (p.sendbuf) { cudaFree(p.sendbuf); p.sendbuf = nullptr; }
    if (p.recvbuf) { cudaFree(p.recvbuf); p.recvbuf = nullptr; }
    p.send_elems = (long long)g.nk * ((long long)(1 + plan->ntr) * p.n_send[0] + p.n_send[1] + p.n_send[2]);
    p.recv_elems = (long long)g.nk * ((long long)(1 + plan->ntr) * p.n_recv[0] + p.n_recv[1] + p.n_recv[2]);
    int rc;
    if (p.send_elems && (rc = dev_alloc(plan, &p.sendbuf, p.send_elems))) return rc;
    if (p.recv_elems && (rc = dev_alloc(plan, &p.recvbuf, p.recv_elems))) return rc;
  }
  plan->peer_ntr = plan->ntr;
  return 0;
}

// One directional sweep over cells [is..ie] x [js..je] (advect_x :329 / advect_y :705),
// all levels with domore_k > 0, all tracers.
static int sweep_impl(mom6adv_plan *plan, int dir, Range rg, double Idt, int scheme, int itt);
static int sweep(mom6adv_plan *plan, int dir, Range rg, double Idt, int scheme, int itt) {
  const long long sp0 = plan->sparse_sweeps;
  const int rc = sweep_impl(plan, dir, rg, Idt, scheme, itt);
  // a dense sweep (any kernel family) may have changed every tracer cell: the neighbours' next pull takes everything
  // (a list-driven sweep marked its tiles itself, k_sparse_scatter)
  if (!rc && plan->dirty && plan->sparse_sweeps == sp0)
    CK(cudaMemsetAsync(plan->dirty + (plan->g.plane >> 5) * plan->g.nk, 1, 1, plan->stream));
  return rc;
}
static int sweep_impl(mom6adv_plan *plan, int dir, Range rg, double Idt, int scheme, int itt) {
  const Geom &g = plan->g;
  cudaStream_t s = plan->stream;
  const double min_h = 0.1 * plan->Angstrom_H;        // :401 / :773
  const double h_neglect = plan->H_subroundoff;       // :403 / :775
  bool any_ad2d = false, any_diag = false;
  for (int m = plan->m_lo; m < plan->m_hi; m++) {
    if (plan->tr[m].diag[dir == 0 ? 2 : 3]) any_ad2d = true;
    for (int d = 0; d < 5; d++) if (plan->tr[m].diag[d]) any_diag = true;
  }
  if (plan->obc_active) {
    if (plan->pull_pending) { CK(cudaStreamWaitEvent(s, plan->ev_pull_done, 0)); plan->pull_pending = false; }
    return sweep_obc(plan, dir, rg, Idt, scheme);
  }
  const bool split_rows = plan->pull_pending && dir == 0 && plan->opt_fast && itt == 1;
  if (plan->pull_pending && !split_rows) {            // any other consumer needs the halo rows now
    CK(cudaStreamWaitEvent(s, plan->ev_pull_done, 0));
    plan->pull_pending = false;
  }
  if (plan->opt_fast) {
    // streaming kernels (fast.cu): mass fields as below, then all tracers in chunks of MOM6ADV_MAXT
    const bool x = (dir == 0);
    int *dom = x ? plan->domore_u : plan->domore_v, *ent = x ? plan->entry_u : plan->entry_v;
    double *res = x ? plan->uhr : plan->vhr;
    // residual iterations: the cells that change are few -- collect them and update in place (sparse.cu).  The list
    // length stays on the device (the kernels stride over it), the staging space is the tracers' idle ping-pong
    // buffers, so the list cannot overflow and the host never waits inside a sweep.  Whether a trip is worth the
    // list-driven kernels is a host-side HINT from the previous trip's counts (either choice gives the same bits).
    const long long dense_above = plan->opt_sp_cap > 0 ? plan->opt_sp_cap : std::max<long long>(4096, field_elems(plan) / 16);
    bool sparse = plan->opt_sparse && itt >= 2 && !(plan->multi && plan->obc_global) && field_elems(plan) < (1LL << 31) &&
                  (plan->est_cells[dir] < 0 || plan->est_cells[dir] <= dense_above);
    plan->est_dense[dir] = !sparse;
    if (sparse && !plan->sp_list) {
      plan->sp_cap = (int)field_elems(plan);
      int rc = dev_alloc(plan, &plan->sp_list, plan->sp_cap); if (rc) return rc;
    }
    {
      ProfScope ps(plan, MOM6ADV_KC_MASS);
      CK(cudaMemsetAsync(plan->sum_dev + 1, 0, 2 * sizeof(int), s));      // [1] cells listed, [2] tiles listed
      launch_flags_begin(s, g, plan->domore_k, dom, ent, x ? rg.js : rg.js - 1, rg.je);
      if (sparse && !plan->tflags) {
        int rc = dev_alloc(plan, &plan->tflags, mass_tile_count(g)); if (rc) return rc;
        rc = dev_alloc(plan, &plan->tlist, mass_tile_count(g)); if (rc) return rc;
        CK(cudaDeviceGetAttribute(&plan->nsm, cudaDevAttrMultiProcessorCount, plan->device));
      }
      if (sparse) {
        // residual trips: a persistent scan of uhr / vhr computes the few active tiles, the update visits only those
        launch_mass_flux_scan(s, dir, g, rg, res, plan->hprev, plan->areaT, min_h, ent, dom, plan->hh, plan->cfl,
                              plan->hc, plan->gq, scheme == SCHEME_PLM, plan->tflags, plan->tlist, plan->sum_dev + 2, plan->nsm);
        launch_mass_update_list(s, dir, g, rg, res, plan->hprev, plan->hh, plan->areaT, x ? plan->uh_neglect : plan->vh_neglect,
                                h_neglect, ent, plan->domore_k, plan->hlst, plan->ihnew, plan->sp_list, plan->sum_dev + 1,
                                plan->sp_cap, plan->tlist, plan->sum_dev + 2, plan->nsm);
        // this trip's list length is the estimate for the next one (read with the convergence word, once per trip)
        CK(cudaMemcpyAsync(plan->sum_dev + 4 + dir, plan->sum_dev + 1, sizeof(int), cudaMemcpyDeviceToDevice, s));
      } else {
        // dense trip of the streaming path: nothing reads CFL itself (k_x_fast / k_y_ring take hc, gq), so it is not written
        CK(cudaMemsetAsync(plan->sum_dev + 4 + dir, 0, sizeof(int), s));
        launch_mass_flux(s, dir, g, rg, res, plan->hprev, plan->areaT, min_h, ent, dom, plan->hh, (any_ad2d || (!x && any_diag)) ? plan->cfl : nullptr,
                         plan->hc, plan->gq, scheme == SCHEME_PLM);      // (CFL itself: only k_ad2d / k_flux_diag read it)
        launch_mass_update(s, dir, g, rg, res, plan->hprev, plan->hh, plan->areaT, x ? plan->uh_neglect : plan->vh_neglect,
                           h_neglect, ent, plan->domore_k, plan->hlst, plan->ihnew, nullptr, nullptr, 0, plan->sum_dev + 4 + dir);
      }
      plan->launches += 3;
    }
    // Tracers with a flux diagnostic of this direction (ad_x | ad_y) or advection_xy go through the DIAG instantiation of
    // the same kernels, as a launch of their own; every other tracer takes the plain one: one associated diagnostic no
    // longer moves the whole registry onto the generic kernels.  ad2d_x / ad2d_y are k-ORDERED sums (:691-699 /
    // :1078-1084): k_ad2d walks k per face column from the sweep's input values -- before the in-place y sweep, after
    // the out-of-place x sweep.
    struct Batch { TrTable tt; DiagTable dg; bool diag; };
    std::vector<Batch> batches;
    // y: the in-place march carries no diagnostics; ad_y / advection_xy of the tracers that have them come from
    // k_flux_diag in front of it (below), so every tracer takes the plain kernel.  (A DIAG instantiation of the march was
    // built and measured: the diagnostic warp's read-modify-write per step is latency-bound while its CTA-mates wait
    // for the ring -- 43.6 ms instead of 31.6 ms per sweep for ad_y on two tracers, profiles/r02_diag_path.md.)
    std::vector<char> in_diag(plan->ntr, 0);
    bool any_dirdiag = false;
    for (int m = plan->m_lo; m < plan->m_hi; m++)
      if (plan->tr[m].diag[x ? 0 : 1] || plan->tr[m].diag[4]) { any_dirdiag = true; if (x || sparse) in_diag[m] = 1; }
    for (int pass = 0; pass < 2; pass++) {
      Batch b; b.tt.n = 0; b.diag = (pass == 1); memset(&b.dg, 0, sizeof b.dg);
      for (int m = plan->m_lo; m < plan->m_hi; m++) {
        const TracerSlot &ts = plan->tr[m];
        const bool d = in_diag[m] != 0;
        if (d != b.diag) continue;
        const int n = b.tt.n++;
        b.tt.src[n] = ts.buf[plan->cur]; b.tt.cu[n] = ts.conc_underflow;
        b.tt.dst[n] = sparse ? ts.buf[1 - plan->cur] : (x ? ts.buf[1 - plan->cur] : ts.buf[plan->cur]);
        b.dg.ad[n] = ts.diag[x ? 0 : 1]; b.dg.axy[n] = ts.diag[4];
        if (b.tt.n == MOM6ADV_MAXT) { batches.push_back(b); b.tt.n = 0; memset(&b.dg, 0, sizeof b.dg); }
      }
      if (b.tt.n > 0) batches.push_back(b);
    }
    const double *mmd = x ? plan->mmx : plan->mmy;
    auto ad2d = [&]() {
      launch_ad2d(s, scheme, dir, g, rg, plan->tr_dev + plan->m_lo, plan->m_hi - plan->m_lo, plan->cur, plan->hh, plan->cfl,
                  plan->ihnew, mmd, ent, plan->domore_k, Idt, sparse ? plan->tflags : nullptr);
      plan->launches++;
    };
    if (sparse) {
      ProfScope pt(plan, MOM6ADV_KC_SPARSE);
      if (any_ad2d) ad2d();      // (reads the live buffers, which the list-driven update changes only afterwards)
      for (const Batch &b : batches) {
        launch_sparse_sweep(s, scheme, dir, g, rg, b.tt, plan->sp_list, plan->sum_dev + 1, plan->nsm, plan->hh, plan->cfl,
                            plan->hlst, plan->ihnew, mmd, ent, plan->dirty, b.diag ? &b.dg : nullptr, plan->IareaT, Idt,
                            plan->tflags);
        plan->launches += 2;
      }
      plan->sparse_sweeps++;
      CK(cudaGetLastError());
      return 0;
    }
    plan->dense_sweeps++;
    ProfScope pt(plan, x ? MOM6ADV_KC_TRACER_X : MOM6ADV_KC_TRACER_Y);
    if (!x && any_ad2d) ad2d();
    if (!x && any_dirdiag) {
      launch_flux_diag(s, scheme, 1, g, rg, plan->tr_dev + plan->m_lo, plan->m_hi - plan->m_lo, plan->cur, plan->hh, plan->cfl,
                       plan->ihnew, plan->mmy, plan->IareaT, ent, plan->domore_k, Idt);
      plan->launches++;
    }
    bool first = true;
    for (const Batch &b : batches) {
      const DiagTable *dg = b.diag ? &b.dg : nullptr;
      if (x && split_rows) {
        // rows of the computational domain first; the halo rows once their tracer values have arrived
        launch_x_fast(s, scheme, g, rg, b.tt, plan->hh, plan->hc, plan->gq, plan->hlst, plan->ihnew, plan->mmx, ent,
                      plan->domore_k, plan->wbase, g.jsc, g.jec, dg, plan->IareaT, Idt);
        plan->launches++;
      } else if (x) {
        launch_x_fast(s, scheme, g, rg, b.tt, plan->hh, plan->hc, plan->gq, plan->hlst, plan->ihnew, plan->mmx, ent,
                      plan->domore_k, plan->wbase, g.jsd, g.jed, dg, plan->IareaT, Idt);
        plan->launches++;
      } else {
        launch_y_fast(s, scheme, g, rg, b.tt, plan->hh, plan->hc, plan->gq, plan->hlst, plan->ihnew, plan->mmy, ent,
                      plan->domore_k);
        plan->launches++;
      }
      first = false;
    }
    (void)first;
    if (x && split_rows) {
      if (plan->pull_pending) { CK(cudaStreamWaitEvent(s, plan->ev_pull_done, 0)); plan->pull_pending = false; }
      for (const Batch &b : batches) {
        const DiagTable *dg = b.diag ? &b.dg : nullptr;
        launch_x_fast(s, scheme, g, rg, b.tt, plan->hh, plan->hc, plan->gq, plan->hlst, plan->ihnew, plan->mmx, ent,
                      plan->domore_k, plan->wbase, g.jsd, g.jsc - 1, dg, plan->IareaT, Idt);
        launch_x_fast(s, scheme, g, rg, b.tt, plan->hh, plan->hc, plan->gq, plan->hlst, plan->ihnew, plan->mmx, ent,
                      plan->domore_k, plan->wbase, g.jec + 1, g.jed, dg, plan->IareaT, Idt);
        plan->launches += 2;
      }
    }
    if (x && any_ad2d) ad2d();             // (plan->cur still names the sweep's input buffers)
    if (x) plan->cur = 1 - plan->cur;      // x is out of place, y in place
    CK(cudaGetLastError());
    return 0;
  }
  if (dir == 0) {
    {
      ProfScope ps(plan, MOM6ADV_KC_MASS);
      launch_flags_begin(s, g, plan->domore_k, plan->domore_u, plan->entry_u, rg.js, rg.je);
      launch_mass_flux(s, 0, g, rg, plan->uhr, plan->hprev, plan->areaT, min_h, plan->entry_u, plan->domore_u, plan->hh, plan->cfl,
                       nullptr, nullptr, 0);
      launch_mass_update(s, 0, g, rg, plan->uhr, plan->hprev, plan->hh, plan->areaT, plan->uh_neglect, h_neglect,
                         plan->entry_u, plan->domore_k, plan->hlst, plan->ihnew, nullptr, nullptr, 0);
    }
    ProfScope pt(plan, MOM6ADV_KC_TRACER_X);
    launch_tracer_sweep(s, scheme, 0, g, rg, plan->tr_dev + plan->m_lo, plan->m_hi - plan->m_lo, plan->cur, plan->hh, plan->cfl, plan->hlst,
                        plan->ihnew, plan->mmx, plan->IareaT, plan->entry_u, plan->domore_k, Idt);
    if (any_ad2d) launch_ad2d(s, scheme, 0, g, rg, plan->tr_dev + plan->m_lo, plan->m_hi - plan->m_lo, plan->cur, plan->hh, plan->cfl, plan->ihnew,
                              plan->mmx, plan->entry_u, plan->domore_k, Idt);
  } else {
    {
      ProfScope ps(plan, MOM6ADV_KC_MASS);
      launch_flags_begin(s, g, plan->domore_k, plan->domore_v, plan->entry_v, rg.js - 1, rg.je);
      launch_mass_flux(s, 1, g, rg, plan->vhr, plan->hprev, plan->areaT, min_h, plan->entry_v, plan->domore_v, plan->hh, plan->cfl,
                       nullptr, nullptr, 0);
      launch_mass_update(s, 1, g, rg, plan->vhr, plan->hprev, plan->hh, plan->areaT, plan->vh_neglect, h_neglect,
                         plan->entry_v, plan->domore_k, plan->hlst, plan->ihnew, nullptr, nullptr, 0);
    }
    ProfScope pt(plan, MOM6ADV_KC_TRACER_Y);
    launch_tracer_sweep(s, scheme, 1, g, rg, plan->tr_dev + plan->m_lo, plan->m_hi - plan->m_lo, plan->cur, plan->hh, plan->cfl, plan->hlst,
                        plan->ihnew, plan->mmy, plan->IareaT, plan->entry_v, plan->domore_k, Idt);
    if (any_ad2d) launch_ad2d(s, scheme, 1, g, rg, plan->tr_dev + plan->m_lo, plan->m_hi - plan->m_lo, plan->cur, plan->hh, plan->cfl, plan->ihnew,
                              plan->mmy, plan->entry_v, plan->domore_k, Idt);
  }
  plan->launches += 4 + (any_ad2d ? 1 : 0);
  plan->cur = 1 - plan->cur;
  CK(cudaGetLastError());
  return 0;
}

// advect_tracer :112-320 on the resident fields.
// The tracers [m_lo, m_hi) are advected; the tracer-independent part (uhr, vhr, hprev, flags) is the same
// whatever the range, which is what lets MIRROR mode pipeline chunks of tracers against the PCIe copies.
// Device tracer table, neighbour mappings and message buffers for the registry as it stands (start of every call
// that makes group passes).
static int prepare_exchange(mom6adv_plan *plan) {
  cudaStream_t s = plan->stream;
  if (plan->multi && plan->comm && plan->opt_pull) {
    // the remap decision is collective: a registry re-bound on some ranks only must not leave the others outside
    // setup_ipc's collectives (one 4-byte all-reduce per call)
    if (!plan->bar_dev) { int rc = dev_alloc(plan, &plan->bar_dev, 4); if (rc) return rc; }
    int need = (plan->tr_dirty || plan->ipc_ntr != plan->ntr) ? 1 : 0;
    CK(cudaMemcpyAsync(plan->bar_dev + 2, &need, sizeof(int), cudaMemcpyHostToDevice, s));
    CKN(ncclAllReduce(plan->bar_dev + 2, plan->bar_dev + 2, 1, ncclInt, ncclMax, plan->comm, s));
    CK(cudaMemcpyAsync(&need, plan->bar_dev + 2, sizeof(int), cudaMemcpyDeviceToHost, s));
    CK(cudaStreamSynchronize(s));
    if (need && plan->ipc_ntr >= 0) plan->ipc_ntr = -2;
  }
  { int rc = sync_tracer_table(plan); if (rc) return rc; }
  { int rc = setup_ipc(plan); if (rc) return rc; }
  { int rc = size_peer_buffers(plan); if (rc) return rc; }
  return 0;
}

static int run_advect(mom6adv_plan *plan, double dt, int x_first_in, int max_iter_in, bool use_vol_prev,
                      bool update_vol_prev, int *n_iter_out, int m_lo, int m_hi) {
  if (n_iter_out) *n_iter_out = 0;
  if (plan->ntr == 0) return MOM6ADV_OK;                                   // :124
  if (m_lo < 0 || m_hi > plan->ntr || m_lo >= m_hi) return fail(plan, MOM6ADV_ERR_INVALID, "advect: bad tracer range");
  for (int m = m_lo; m < m_hi; m++) {
    if (!plan->tr[m].buf[0]) return fail(plan, MOM6ADV_ERR_NOT_INIT, "advect: tracer storage missing");
    if (plan->tr[m].cur != plan->tr[m_lo].cur) return fail(plan, MOM6ADV_ERR_INVALID, "advect: tracer buffers out of step");
  }
  plan->m_lo = m_lo; plan->m_hi = m_hi; plan->cur = plan->tr[m_lo].cur;
  const Geom &g = plan->g;
  cudaStream_t s = plan->stream;
  CK(cudaSetDevice(plan->device));
  { int rc = prepare_exchange(plan); if (rc) return rc; }
  const int is = g.isc, ie = g.iec, js = g.jsc, je = g.jec;
  int stencil = 2;                                                          // :118
  const bool use_PPM_stencil = plan->cfg.usePPM && !plan->cfg.useHuynhStencilBug;   // :129
  if (use_PPM_stencil) stencil = 3;                                         // :130
  const int scheme = !plan->cfg.usePPM ? SCHEME_PLM : (plan->cfg.useHuynh ? SCHEME_PPM_H3 : SCHEME_PPM_CW);
  const int halo_min = std::min(std::min(is - g.isd, g.ied - ie), std::min(js - g.jsd, g.jed - je));
  if (halo_min < stencil)
    return fail(plan, MOM6ADV_ERR_UNSUPPORTED, "advect: halo narrower than the advection stencil");
  if (!(dt > 0.0) || !(plan->cfg.dt_baroclinic > 0.0)) return fail(plan, MOM6ADV_ERR_INVALID, "advect: dt must be > 0");
  bool x_first = ((plan->cfg.first_direction % 2) == 0);                    // :126
  const double Idt = 1.0 / dt;                                              // :133
  int max_iter = 2 * (int)std::ceil(dt / plan->cfg.dt_baroclinic) + 1;      // :135
  if (max_iter_in >= 0) max_iter = max_iter_in;                             // :137
  if (x_first_in >= 0) x_first = (x_first_in != 0);                         // :138
  if (use_vol_prev && !plan->vol_prev) return fail(plan, MOM6ADV_ERR_INVALID, "advect: vol_prev was not uploaded");

  CK(cudaEventRecord(plan->ev0, s));
  // :112-113, :156
  CK(cudaMemsetAsync(plan->domore_u, 0, sizeof(int) * g.R * g.nk, s));
  CK(cudaMemsetAsync(plan->domore_v, 0, sizeof(int) * g.R * g.nk, s));
  { ProfScope ps(plan, MOM6ADV_KC_PROLOGUE);
  launch_init_residuals(s, g, plan->h_end, plan->uhtr, plan->vhtr, use_vol_prev ? plan->vol_prev : nullptr,
                        plan->areaT, plan->uhr, plan->vhr, plan->hprev);    // :152-178
  }
  launch_fill_int(s, plan->domore_k, g.nk, 1);                               // :156
  plan->launches += 2;
  for (int m = m_lo; m < m_hi; m++)                                         // :192-198
    for (int d = 0; d < 5; d++)
      if (plan->tr[m].diag[d])
        CK(cudaMemsetAsync(plan->tr[m].diag[d], 0, sizeof(double) * ((d == 2 || d == 3) ? g.plane : field_elems(plan)), s));

  plan->est_cells[0] = plan->est_cells[1] = -1;
  int isv = is, iev = ie, jsv = js, jev = je;                               // :201
  int itt;
  for (itt = 1; itt <= max_iter; itt++) {                                   // :203
    if (isv > is - stencil) {                                               // :205
      bool defer = false;
      // (deferred only when the convergence all-reduce of this trip is certain to follow -- nsten_halo == 1 and more
      // than one trip allowed: it is what orders the neighbours' deferred reads before anybody's next writes)
      if (itt == 1 && x_first && plan->opt_overlap && plan->opt_fast && !plan->obc_active && plan->dom.npi == 1 && halo_min / stencil == 1 && max_iter > 1) {
        defer = true;
        for (int m = m_lo; m < m_hi; m++) for (int d = 0; d < 5; d++) if (plan->tr[m].diag[d]) defer = false;
      }
      { int rc = group_pass(plan, itt == 1, defer); if (rc) return rc; }     // :206
      const int nsten_halo = halo_min / stencil;                            // :208
      isv = is - nsten_halo * stencil; jsv = js - nsten_halo * stencil;     // :209
      iev = ie + nsten_halo * stencil; jev = je + nsten_halo * stencil;     // :210
      if ((nsten_halo > 1) || (itt == 1)) {                                 // :213
        launch_rescan(s, g, plan->uhr, plan->vhr, plan->domore_k, plan->domore_u, plan->domore_v, isv, iev, jsv, jev, stencil);
        launch_domore_k(s, g, plan->domore_k, plan->domore_u, plan->domore_v, jsv, jev, jsv + stencil - 1, jev - stencil);
        plan->launches += 2;
      }
    }
    isv += stencil; iev -= stencil; jsv += stencil; jev -= stencil;         // :238-239
    int rc;
    if (x_first) {                                                          // :249-269
      Range rx = {isv, iev, jsv - stencil, jev + stencil};
      Range ry = {isv, iev, jsv, jev};
      if ((rc = sweep(plan, 0, rx, Idt, scheme, itt))) return rc;
      if ((rc = sweep(plan, 1, ry, Idt, scheme, itt))) return rc;
      launch_domore_k(s, g, plan->domore_k, plan->domore_u, plan->domore_v, jsv - stencil, jev + stencil, jsv - 1, jev);
    } else {                                                                // :271-290
      Range ry = {isv - stencil, iev + stencil, jsv, jev};
      Range rx = {isv, iev, jsv, jev};
      if ((rc = sweep(plan, 1, ry, Idt, scheme, itt))) return rc;
      if ((rc = sweep(plan, 0, rx, Idt, scheme, itt))) return rc;
      launch_domore_k(s, g, plan->domore_k, plan->domore_u, plan->domore_v, jsv, jev, jsv - 1, jev);
    }
    plan->launches += 1;
    if (itt >= max_iter) break;                                             // :297-299
    const bool test = (isv > is - stencil);                                 // :302
    // ONE host readback per trip: the convergence word and, with it, the two counts that hint whether the next trip's
    // sweeps are worth the list-driven kernels.  (PLM with halo 4 has no test after the first trip: read the counts alone.)
    const bool want_est = plan->opt_fast && plan->opt_sparse && !plan->obc_active && (test || itt == 1);
    if (test) {
      if (plan->multi && plan->comm)                                        // :305 sum_across_PEs(domore_k, nz)
        CKN(ncclAllReduce(plan->domore_k, plan->domore_k, g.nk, ncclInt, ncclSum, plan->comm, s));
      launch_sum_domore_k(s, plan->domore_k, g.nk, plan->sum_dev);          // :307 do_any
      plan->launches += 1;
    }
    // multi-block: the list-driven trip leaves the live buffer where it is and the dense one flips it, so every rank
    // must take the same kind of trip -- the hint becomes the maximum over the ranks
    if (plan->multi && plan->comm && (test || itt == 1))
      CKN(ncclAllReduce(plan->sum_dev + 4, plan->sum_dev + 4, 2, ncclInt, ncclMax, plan->comm, s));
    if (test || want_est) {
      launch_publish(s, plan->sum_host_dev, plan->sum_dev, 6);
      CK(cudaStreamSynchronize(s));
      for (int d = 0; d < 2; d++) plan->est_cells[d] = plan->est_dense[d] ? 2LL * plan->sum_host[4 + d] : plan->sum_host[4 + d];
      if (test && plan->sum_host[0] == 0) break;                            // :308-310
    }
  }
  if (plan->pull_pending) { CK(cudaStreamWaitEvent(s, plan->ev_pull_done, 0)); plan->pull_pending = false; }
  if (n_iter_out) *n_iter_out = std::min(itt, max_iter);
  if (update_vol_prev && use_vol_prev)                                      // :318-320
    CK(cudaMemcpyAsync(plan->vol_prev, plan->hprev, sizeof(double) * field_elems(plan), cudaMemcpyDeviceToDevice, s));
  CK(cudaEventRecord(plan->ev1, s));
  CK(cudaStreamSynchronize(s));
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, plan->ev0, plan->ev1));
  plan->last_ms = ms;
  for (int m = m_lo; m < m_hi; m++) plan->tr[m].cur = plan->cur;
  prof_collect(plan);
  CK(cudaGetLastError());
  return MOM6ADV_OK;
}

// ---------------------------------------------------------------------------
// RESIDENT-mode data movement
// ---------------------------------------------------------------------------
static int field_info(mom6adv_plan *plan, int field, double ***pp, int *ilo, int *jlo) {
  const Geom &g = plan->g;
  *ilo = g.isd; *jlo = g.jsd;
  switch (field) {
    case MOM6ADV_F_H_END: *pp = &plan->h_end; break;
    case MOM6ADV_F_UHTR: *pp = &plan->uhtr; *ilo = g.IsdB; break;
    case MOM6ADV_F_VHTR: *pp = &plan->vhtr; *jlo = g.JsdB; break;
    case MOM6ADV_F_VOL_PREV: *pp = &plan->vol_prev; break;
    case MOM6ADV_F_UHR: *pp = &plan->uhr; *ilo = g.IsdB; break;
    case MOM6ADV_F_VHR: *pp = &plan->vhr; *jlo = g.JsdB; break;
    case MOM6ADV_F_HPREV: *pp = &plan->hprev; break;
    default: return fail(plan, MOM6ADV_ERR_INVALID, "unknown field id");
  }
  return 0;
}

extern "C" int mom6adv_upload_field(mom6adv_plan *plan, int field, const double *src) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "upload_field: plan not initialised");
  if (!src) return fail(plan, MOM6ADV_ERR_INVALID, "upload_field: NULL source");
  CK(cudaSetDevice(plan->device));
  double **pp; int ilo, jlo;
  int rc = field_info(plan, field, &pp, &ilo, &jlo);
  if (rc) return rc;
  if (!*pp) { rc = dev_alloc(plan, pp, field_elems(plan)); if (rc) return rc; }
  rc = upload(plan, *pp, src, ilo, jlo, plan->g.nk);
  if (rc) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

extern "C" int mom6adv_download_field(mom6adv_plan *plan, int field, double *dst) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "download_field: plan not initialised");
  if (!dst) return fail(plan, MOM6ADV_ERR_INVALID, "download_field: NULL destination");
  CK(cudaSetDevice(plan->device));
  double **pp; int ilo, jlo;
  int rc = field_info(plan, field, &pp, &ilo, &jlo);
  if (rc) return rc;
  if (!*pp) return fail(plan, MOM6ADV_ERR_INVALID, "download_field: field not resident");
  rc = download(plan, dst, *pp, ilo, jlo, plan->g.nk);
  if (rc) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

extern "C" int mom6adv_upload_tracer(mom6adv_plan *plan, int m, const double *src) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "upload_tracer: plan not initialised");
  if (m < 0 || m >= plan->ntr) return fail(plan, MOM6ADV_ERR_INVALID, "upload_tracer: tracer index out of range");
  if (!src) return fail(plan, MOM6ADV_ERR_INVALID, "upload_tracer: NULL source");
  CK(cudaSetDevice(plan->device));
  int rc = upload(plan, plan->tr[m].buf[plan->tr[m].cur], src, plan->g.isd, plan->g.jsd, plan->g.nk);
  if (rc) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

extern "C" int mom6adv_download_tracer(mom6adv_plan *plan, int m, double *dst) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "download_tracer: plan not initialised");
  if (m < 0 || m >= plan->ntr) return fail(plan, MOM6ADV_ERR_INVALID, "download_tracer: tracer index out of range");
  if (!dst) return fail(plan, MOM6ADV_ERR_INVALID, "download_tracer: NULL destination");
  CK(cudaSetDevice(plan->device));
  int rc = download(plan, dst, plan->tr[m].buf[plan->tr[m].cur], plan->g.isd, plan->g.jsd, plan->g.nk);
  if (rc) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

static void diag_info(const Geom &g, int kind, int *ilo, int *jlo, int *nlev) {
  *ilo = g.isd; *jlo = g.jsd; *nlev = g.nk;
  if (kind == 0 || kind == 2 || kind == 5 || kind == 7) *ilo = g.IsdB;      // 5..8: df_x, df_y, df2d_x, df2d_y
  if (kind == 1 || kind == 3 || kind == 6 || kind == 8) *jlo = g.JsdB;
  if (kind == 2 || kind == 3 || kind == 7 || kind == 8) *nlev = 1;
}

extern "C" int mom6adv_download_diag(mom6adv_plan *plan, int m, int kind, double *dst) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "download_diag: plan not initialised");
  if (m < 0 || m >= plan->ntr || kind < 0 || kind > 8) return fail(plan, MOM6ADV_ERR_INVALID, "download_diag: bad index");
  const double *field = kind <= 4 ? plan->tr[m].diag[kind] : plan->tr[m].df[kind - 5];
  if (!field) return fail(plan, MOM6ADV_ERR_INVALID, "download_diag: diagnostic not associated");
  CK(cudaSetDevice(plan->device));
  int ilo, jlo, nlev;
  diag_info(plan->g, kind, &ilo, &jlo, &nlev);
  int rc = download(plan, dst, field, ilo, jlo, nlev);
  if (rc) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

// ---------------------------------------------------------------------------
// Either side of advect_tracer: with these on the device the accumulated transports never
// exist on the host and a column diagnostic costs one 2-D field of D2H.  Scratch: the mass fields of the sweeps
// (hh, cfl, hc), which are dead between two advect calls.
// ---------------------------------------------------------------------------
extern "C" int mom6adv_accumulate_transports(mom6adv_plan *plan, const double *uh, const double *vh, double dt) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "accumulate_transports: plan not initialised");
  if (!uh || !vh) return fail(plan, MOM6ADV_ERR_INVALID, "accumulate_transports: NULL transport");
  CK(cudaSetDevice(plan->device));
  const Geom &g = plan->g;
  // MOM_hor_index.F90:87-96: IscB = isc-1 in symmetric memory (IsdB = isd-1), isc otherwise; Isq = IscB (RK2:409)
  const int IscB = (g.IsdB == g.isd - 1) ? g.isc - 1 : g.isc, JscB = (g.JsdB == g.jsd - 1) ? g.jsc - 1 : g.jsc;
  // MOM_dynamics_split_RK2.F90:1045-1052: j=js-2,je+2 ; I=Isq-2,Ieq+2   and   J=Jsq-2,Jeq+2 ; i=is-2,ie+2
  // (clipped to the data domain: the reference's declared bounds)
  int rc = upload(plan, plan->hh, uh, g.IsdB, g.jsd, g.nk);
  if (rc) return rc;
  launch_accumulate(plan->stream, g, plan->uhtr, plan->hh, dt, std::max(IscB - 2, g.IsdB), std::min(g.iec + 2, g.ied),
                    std::max(g.jsc - 2, g.jsd), std::min(g.jec + 2, g.jed));
  rc = upload(plan, plan->cfl, vh, g.isd, g.JsdB, g.nk);     // (a second scratch: the first copy may still be in flight)
  if (rc) return rc;
  launch_accumulate(plan->stream, g, plan->vhtr, plan->cfl, dt, std::max(g.isc - 2, g.isd), std::min(g.iec + 2, g.ied),
                    std::max(JscB - 2, g.JsdB), std::min(g.jec + 2, g.jed));
  plan->launches += 2;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

extern "C" int mom6adv_zero_transports(mom6adv_plan *plan) {      // MOM.F90:1465-1466
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "zero_transports: plan not initialised");
  CK(cudaSetDevice(plan->device));
  CK(cudaMemsetAsync(plan->uhtr, 0, sizeof(double) * field_elems(plan), plan->stream));
  CK(cudaMemsetAsync(plan->vhtr, 0, sizeof(double) * field_elems(plan), plan->stream));
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

extern "C" int mom6adv_column_diag(mom6adv_plan *plan, int m, int kind, const double *h_diag, double H_to_m,
                                   double *out2d) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "column_diag: plan not initialised");
  if (m < 0 || m >= plan->ntr || kind < 0 || kind > 2) return fail(plan, MOM6ADV_ERR_INVALID, "column_diag: bad index");
  if (!out2d) return fail(plan, MOM6ADV_ERR_INVALID, "column_diag: NULL destination");
  CK(cudaSetDevice(plan->device));
  const Geom &g = plan->g;
  const TracerSlot &ts = plan->tr[m];
  double *out = plan->hc;                                  // one plane of scratch
  if (kind == MOM6ADV_CD_ADV_XY_2D) {
    if (!ts.diag[4]) return fail(plan, MOM6ADV_ERR_INVALID, "column_diag: advection_xy is not associated");
    launch_col_sum(plan->stream, g, ts.diag[4], out);     // MOM_tracer_registry.F90:786-791
  } else {
    if (!h_diag) return fail(plan, MOM6ADV_ERR_INVALID, "column_diag: NULL h_diag");
    if (!ts.buf[ts.cur]) return fail(plan, MOM6ADV_ERR_INVALID, "column_diag: tracer is not resident");
    int rc = upload(plan, plan->cfl, h_diag, g.isd, g.jsd, g.nk);
    if (rc) return rc;
    if (kind == MOM6ADV_CD_ZINT) launch_col_zint(plan->stream, g, plan->cfl, ts.buf[ts.cur], H_to_m, out);      // :796-802
    else launch_col_zint100(plan->stream, g, plan->cfl, ts.buf[ts.cur], H_to_m, plan->sum_dev + 3, out);        // :745-766,805-811
  }
  plan->launches += (kind == MOM6ADV_CD_ZINT_100M) ? 2 : 1;
  CK(cudaGetLastError());
  int rc = download(plan, out2d, out, g.isd, g.jsd, 1);
  if (rc) return rc;
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

// ---------------------------------------------------------------------------
// tracer_hordiff's diffusion iteration (MOM_tracer_hor_diff.F90:539-607) on the resident tracers; kernels in hordiff.cu.
// Scratch: Coef_x -> hh, Coef_y -> cfl, Ihdxdy -> hlst (advect_tracer's per-sweep work fields, free between calls).
// ---------------------------------------------------------------------------
static int hordiff_run(mom6adv_plan *plan, const double *h, const double *khdt_x, const double *khdt_y,
                       int num_itts, double h_neglect, int Diffuse_ML_interior, int nkml, int nk_rho_varies,
                       double ML_KhTr_scale, bool with_df, double Idt);

extern "C" int mom6adv_hordiff(mom6adv_plan *plan, const double *h, const double *khdt_x, const double *khdt_y,
                               int num_itts, double h_neglect, int Diffuse_ML_interior, int nkml, int nk_rho_varies,
                               double ML_KhTr_scale, int want_df) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "hordiff: plan not initialised");
  if (want_df) return fail(plan, MOM6ADV_ERR_UNSUPPORTED, "hordiff: this entry point does not produce the df_x / df_y / df2d_x / df2d_y diagnostics (mom6adv_hordiff_diag does, behind the option hordiff_df)");
  return hordiff_run(plan, h, khdt_x, khdt_y, num_itts, h_neglect, Diffuse_ML_interior, nkml, nk_rho_varies, ML_KhTr_scale,
                     false, 0.0);
}

// The same call with the diffusive-flux diagnostics (MOM_tracer_hor_diff.F90:388-405 zeroing, :576-591 accumulation) of
// the tracers that have them bound (mom6adv_bind_tracer_df); dt is tracer_hordiff's dt (Idt = 1/dt, :204).
// NOT YET RUN ON A GPU (the round's GPU time was spent): refused unless the option "hordiff_df" is set.
extern "C" int mom6adv_hordiff_diag(mom6adv_plan *plan, const double *h, const double *khdt_x, const double *khdt_y,
                                    int num_itts, double h_neglect, int Diffuse_ML_interior, int nkml, int nk_rho_varies,
                                    double ML_KhTr_scale, double dt) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "hordiff_diag: plan not initialised");
  if (!plan->opt_hordiff_df)
    return fail(plan, MOM6ADV_ERR_UNSUPPORTED, "hordiff_diag: the df_x / df_y / df2d_x / df2d_y path has not been verified on a GPU yet; set the option hordiff_df to use it");
  if (!(dt > 0.0)) return fail(plan, MOM6ADV_ERR_INVALID, "hordiff_diag: dt must be > 0");
  return hordiff_run(plan, h, khdt_x, khdt_y, num_itts, h_neglect, Diffuse_ML_interior, nkml, nk_rho_varies, ML_KhTr_scale,
                     true, 1.0 / dt);
}

static int hordiff_run(mom6adv_plan *plan, const double *h, const double *khdt_x, const double *khdt_y,
                       int num_itts, double h_neglect, int Diffuse_ML_interior, int nkml, int nk_rho_varies,
                       double ML_KhTr_scale, bool with_df, double Idt) {
  if (!khdt_x || !khdt_y) return fail(plan, MOM6ADV_ERR_INVALID, "hordiff: NULL khdt_x / khdt_y");
  if (num_itts < 1) return fail(plan, MOM6ADV_ERR_INVALID, "hordiff: num_itts must be >= 1");
  if (plan->ntr == 0) return MOM6ADV_OK;
  const Geom &g = plan->g;
  for (int m = 0; m < plan->ntr; m++) {
    if (!plan->tr[m].buf[0]) return fail(plan, MOM6ADV_ERR_NOT_INIT, "hordiff: tracer storage missing");
    if (plan->tr[m].cur != plan->tr[0].cur) return fail(plan, MOM6ADV_ERR_INVALID, "hordiff: tracer buffers out of step");
  }
  if (std::min(std::min(g.isc - g.isd, g.ied - g.iec), std::min(g.jsc - g.jsd, g.jed - g.jec)) < 1)
    return fail(plan, MOM6ADV_ERR_UNSUPPORTED, "hordiff: needs a halo of at least one point");
  CK(cudaSetDevice(plan->device));
  cudaStream_t s = plan->stream;
  plan->m_lo = 0; plan->m_hi = plan->ntr; plan->cur = plan->tr[0].cur;
  { int rc = prepare_exchange(plan); if (rc) return rc; }
  if (!plan->khdt_x) {
    int rc = dev_alloc(plan, &plan->khdt_x, g.plane); if (rc) return rc;
    rc = dev_alloc(plan, &plan->khdt_y, g.plane); if (rc) return rc;
  }
  int rc;
  if (h) { if ((rc = upload(plan, plan->h_end, h, g.isd, g.jsd, g.nk))) return rc; CK(cudaStreamSynchronize(s)); }
  if ((rc = upload(plan, plan->khdt_x, khdt_x, g.IsdB, g.jsd, 1))) return rc;
  CK(cudaStreamSynchronize(s));              // the staging buffer is reused by the next upload
  if ((rc = upload(plan, plan->khdt_y, khdt_y, g.isd, g.JsdB, 1))) return rc;

  HdParams hp;
  hp.I_numitts = 1.0 / ((double)num_itts);                                  // :380 / :384 / :386
  hp.ml_scale = ML_KhTr_scale; hp.h_neglect = h_neglect;
  hp.diffuse_ml = Diffuse_ML_interior ? 1 : 0; hp.nkml = nkml; hp.nk_rho_varies = nk_rho_varies;
  CK(cudaEventRecord(plan->ev0, s));
  if (with_df)                                                              // :388-405
    for (int m = 0; m < plan->ntr; m++) for (int d = 0; d < 4; d++) if (plan->tr[m].df[d])
      CK(cudaMemsetAsync(plan->tr[m].df[d], 0, sizeof(double) * (size_t)(d >= 2 ? g.plane : field_elems(plan)), s));
  launch_hd_coef(s, g, hp, plan->h_end, plan->khdt_x, plan->khdt_y, plan->IareaT, plan->hh, plan->cfl, plan->hlst);   // :552-566
  plan->launches++;
  for (int itt = 1; itt <= num_itts; itt++) {                               // :539
    // every rank's previous iteration must be complete before its rows are read: the opening fence each time
    { int rc2 = group_pass(plan, true, false, PASS_TRACERS); if (rc2) return rc2; } // :540
    const int src = plan->cur;
    for (int m0 = 0; m0 < plan->ntr; m0 += MOM6ADV_MAXT) {
      TrTable tt;
      tt.n = std::min(MOM6ADV_MAXT, plan->ntr - m0);
      for (int q = 0; q < tt.n; q++) {
        const TracerSlot &ts = plan->tr[m0 + q];
        tt.src[q] = ts.buf[src]; tt.dst[q] = ts.buf[1 - src]; tt.cu[q] = ts.conc_underflow;
      }
      if (with_df) {                                                        // :576-591, from the values before the update
        DfTable df; bool any3 = false, any2 = false;
        for (int q = 0; q < MOM6ADV_MAXT; q++) {
          const TracerSlot *ts = q < tt.n ? &plan->tr[m0 + q] : nullptr;
          df.dfx[q] = ts ? ts->df[0] : nullptr; df.dfy[q] = ts ? ts->df[1] : nullptr;
          df.d2x[q] = ts ? ts->df[2] : nullptr; df.d2y[q] = ts ? ts->df[3] : nullptr;
          any3 = any3 || df.dfx[q] || df.dfy[q]; any2 = any2 || df.d2x[q] || df.d2y[q];
        }
        if (any3 || any2) {
          launch_hd_df(s, g, hp, tt, df, any3, any2, plan->hh, plan->cfl, Idt);
          plan->launches += (any3 ? 1 : 0) + (any2 ? 1 : 0);
        }
      }
      launch_hd_apply(s, g, hp, tt, plan->hh, plan->cfl, plan->hlst);       // :568-605
      plan->launches++;
    }
    CK(cudaGetLastError());
    plan->cur = 1 - src;
    for (int m = 0; m < plan->ntr; m++) plan->tr[m].cur = plan->cur;
  }
  if (plan->dirty) CK(cudaMemsetAsync(plan->dirty + (g.plane >> 5) * g.nk, 1, 1, s));   // every tracer cell may have changed
  CK(cudaEventRecord(plan->ev1, s));
  CK(cudaStreamSynchronize(s));
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, plan->ev0, plan->ev1));
  plan->last_ms = ms;
  return MOM6ADV_OK;
}

extern "C" int mom6adv_advect_resident(mom6adv_plan *plan, double dt, int x_first, int max_iter_in,
                                       int use_vol_prev, int update_vol_prev, int *n_iter) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "advect: tracer_advect_init must be called before advect_tracer");
  return run_advect(plan, dt, x_first, max_iter_in, use_vol_prev != 0, update_vol_prev > 0, n_iter, 0, plan->ntr);
}

// ---------------------------------------------------------------------------
// The iteration loop of offline_advection_ale (MOM_offline_main.F90:254-312, 320-411, 414-415) on the resident fields;
// element-wise kernels and the reproducing sum in offline.cu.  Field roles: see the head of offline.cu.
// ---------------------------------------------------------------------------
static int offline_residual(mom6adv_plan *plan, const double *u, const double *v, double HL2_to_kg_scale, double scale_out,
                            double *out) {
  cudaStream_t s = plan->stream;
  CK(cudaMemsetAsync(plan->off_acc, 0, sizeof(unsigned long long) * 6, s));
  launch_off_residual(s, plan->g, u, v, plan->h_end, plan->areaT, plan->H_subroundoff, HL2_to_kg_scale, plan->off_acc);
  plan->launches++;
  if (plan->multi && plan->comm)                                             // sum_across_PEs of the integers, MOM_coms.F90:195-197
    CKN(ncclAllReduce(plan->off_acc, plan->off_acc, 6, ncclUint64, ncclSum, plan->comm, s));
  long long ints[6];
  CK(cudaMemcpyAsync(ints, plan->off_acc, sizeof ints, cudaMemcpyDeviceToHost, s));
  CK(cudaStreamSynchronize(s));
  // regularize_ints, MOM_coms.F90:653-691
  const long long prec = 1LL << 46;
  const double r_prec = 70368744177664.0, I_prec = 1.0 / 70368744177664.0;
  for (int i = 5; i >= 1; i--) if (std::llabs(ints[i]) >= prec) {
    const int num_carry = (int)(ints[i] * I_prec);
    ints[i] = ints[i] - num_carry * prec;
    ints[i - 1] = ints[i - 1] + num_carry;
  }
  bool positive = true;
  for (int i = 0; i < 6; i++) if (std::llabs(ints[i]) > 0) { if (ints[i] < 0) positive = false; break; }
  if (positive) { for (int i = 5; i >= 1; i--) if (ints[i] < 0) { ints[i] += prec; ints[i - 1] -= 1; } }
  else { for (int i = 5; i >= 1; i--) if (ints[i] > 0) { ints[i] -= prec; ints[i - 1] += 1; } }
  const double pr[6] = {r_prec * r_prec, r_prec, 1.0, 1.0 / r_prec, 1.0 / (r_prec * r_prec), 1.0 / (r_prec * r_prec * r_prec)};
  double r = 0.0;                                                            // ints_to_real :550-559
  for (int i = 0; i < 6; i++) r = r + pr[i] * (double)ints[i];
  *out = scale_out * r;                                                      // MOM_offline_main.F90:647-648
  return 0;
}

extern "C" int mom6adv_offline_advect(mom6adv_plan *plan, double dt_offline, int num_off_iter, int off_ale_mod,
                                      int x_before_y_in, double min_residual, double H_to_kg_m2, double L_to_m,
                                      double kg_m2_to_H, double m_to_L, int *converged_out, int *n_iter_out,
                                      double *tot_residual_out) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "offline_advect: plan not initialised");
  if (converged_out) *converged_out = 0;
  if (n_iter_out) *n_iter_out = 0;
  if (num_off_iter < 1) return fail(plan, MOM6ADV_ERR_INVALID, "offline_advect: num_off_iter must be >= 1");
  for (int iter = 1; iter <= num_off_iter; iter++)
    if (off_ale_mod <= 0 || iter % off_ale_mod == 0)
      return fail(plan, MOM6ADV_ERR_UNSUPPORTED, "offline_advect: the ALE step inside the loop (MODULO(iter, off_ale_mod) == 0) is not built; keep the Fortran loop for such runs");
  if (plan->ntr == 0) return fail(plan, MOM6ADV_ERR_INVALID, "offline_advect: no tracers registered");
  if (plan->multi && !plan->comm) return fail(plan, MOM6ADV_ERR_COMM, "offline_advect: mom6adv_comm_init has not been called on this multi-block plan");
  const Geom &g = plan->g;
  CK(cudaSetDevice(plan->device));
  cudaStream_t s = plan->stream;
  int rc;
  if (!plan->vol_prev && (rc = dev_alloc(plan, &plan->vol_prev, field_elems(plan)))) return rc;
  if (!plan->off_acc) CK(cudaMalloc((void **)&plan->off_acc, sizeof(unsigned long long) * 6));
  const size_t fb = sizeof(double) * (size_t)field_elems(plan);
  if (!plan->ev_off0) { CK(cudaEventCreate(&plan->ev_off0)); CK(cudaEventCreate(&plan->ev_off1)); }
  cudaEvent_t e0 = plan->ev_off0, e1 = plan->ev_off1;
  CK(cudaEventRecord(e0, s));
  plan->m_lo = 0; plan->m_hi = plan->ntr; plan->cur = plan->tr[0].cur;
  if ((rc = prepare_exchange(plan))) return rc;
  // symmetric memory: the boundary faces is-1 / js-1 have an owner unless they lie on a closed edge of the global domain
  const bool symmetric = (g.IsdB == g.isd - 1);
  const bool keep_w = symmetric && !plan->dom.cyclic_x && (g.isc - 1 + plan->dom.ioff < 1);
  const bool keep_s = symmetric && !plan->dom.cyclic_y && (g.jsc - 1 + plan->dom.joff < 1);
  bool x_before_y = x_before_y_in != 0;                                      // :254
  const double HL2_to_kg_scale = H_to_kg_m2 * (L_to_m * L_to_m);             // :636
  const double rts_scale = 0.5 * kg_m2_to_H * (m_to_L * m_to_L);             // :647
  launch_off_init(s, g, plan->h_end);                                        // :270, :304-306 (uhtr_sub = uhtr, vhtr_sub = vhtr: the same fields)
  plan->launches++;
  double tot_residual = 0.0, prev_tot_residual = std::numeric_limits<double>::quiet_NaN();   // (unset before :409)
  if ((rc = offline_residual(plan, plan->uhtr, plan->vhtr, HL2_to_kg_scale, rts_scale, &tot_residual))) return rc;   // :312
  int nrec = 0;
  if (tot_residual_out) tot_residual_out[nrec] = tot_residual;
  nrec++;
  bool converged = false;                                                    // :276
  int iter;
  for (iter = 1; iter <= num_off_iter; iter++) {                             // :320
    launch_off_vol(s, g, plan->h_end, plan->areaT, plan->vol_prev);          // :322-325
    plan->launches++;
    if ((rc = run_advect(plan, dt_offline, x_before_y ? 1 : 0, 1, true, true, nullptr, 0, plan->ntr))) return rc;   // :334-336
    x_before_y = !x_before_y;                                                // :339
    launch_off_post(s, g, plan->vol_prev, plan->IareaT, plan->h_end, plan->uhr, plan->vhr, plan->uhtr, plan->vhtr);   // :342-344, :382-385
    // :386-387 pass_var(h_new) ; pass_vector(uhtr_sub, vhtr_sub) through the plan's group pass, which moves hprev, uhr
    // and vhr: h_new rides in hprev; uhr / vhr ARE uhtr / vhtr, equal to uhtr_sub / vhtr_sub on the computational
    // faces, and their halo faces take the owner's value of the same face (the bits this rank computed itself).  The
    // halos then go where the reference has them: h_new's, uhtr_sub's, vhtr_sub's.
    CK(cudaMemcpyAsync(plan->hprev, plan->h_end, fb, cudaMemcpyDeviceToDevice, s));
    if ((rc = group_pass(plan, true, false, PASS_MASS))) return rc;
    launch_off_halos(s, g, keep_w, keep_s, plan->hprev, plan->uhr, plan->vhr, plan->h_end, plan->uhtr, plan->vhtr);
    plan->launches += 2;
    if ((rc = offline_residual(plan, plan->uhr, plan->vhr, HL2_to_kg_scale, rts_scale, &tot_residual))) return rc;   // :391
    if (tot_residual_out) tot_residual_out[nrec] = tot_residual;
    nrec++;
    if (tot_residual == 0.0) { converged = true; break; }                    // :397-402
    if ((tot_residual == prev_tot_residual) || (tot_residual < min_residual)) { converged = false; break; }   // :404-407
    prev_tot_residual = tot_residual;                                        // :409
  }
  // :414 h_pre is h_new already ; :415 pass_vector(uhtr, vhtr): uhr / vhr were passed in place after the last trip
  CK(cudaEventRecord(e1, s));
  CK(cudaStreamSynchronize(s));
  float ms = 0.f;
  CK(cudaEventElapsedTime(&ms, e0, e1));
  plan->last_ms = ms;
  CK(cudaGetLastError());
  if (converged_out) *converged_out = converged ? 1 : 0;
  if (n_iter_out) *n_iter_out = iter > num_off_iter ? num_off_iter : iter;
  return MOM6ADV_OK;
}

// ---------------------------------------------------------------------------
// MIRROR mode: one call == one reference advect_tracer call on host arrays.
// ---------------------------------------------------------------------------
// Strided DMA between a Fortran-layout host array (ilo:ied, jlo:jed, nlev) and the padded HBM field:
// no staging buffer and no repack kernel, so copy engines and SMs work at the same time.
static int xfer3d(mom6adv_plan *plan, cudaStream_t st, double *dev, const double *host, int ilo, int jlo, int nlev,
                  bool to_dev) {
  const Geom &g = plan->g;
  const size_t ni = (size_t)(g.ied - ilo + 1), nj = (size_t)(g.jed - jlo + 1);
  cudaMemcpy3DParms p;
  memset(&p, 0, sizeof p);
  const cudaPitchedPtr hp = make_cudaPitchedPtr((void *)host, ni * 8, ni * 8, nj);
  const cudaPitchedPtr dp = make_cudaPitchedPtr((void *)dev, (size_t)g.P * 8, (size_t)g.P * 8, (size_t)g.R);
  const cudaPos dpos = make_cudaPos((size_t)(ilo + g.coff) * 8, (size_t)(jlo + g.roff), 0);
  if (to_dev) { p.srcPtr = hp; p.dstPtr = dp; p.dstPos = dpos; p.kind = cudaMemcpyHostToDevice; }
  else { p.srcPtr = dp; p.srcPos = dpos; p.dstPtr = hp; p.kind = cudaMemcpyDeviceToHost; }
  p.extent = make_cudaExtent(ni * 8, nj, (size_t)nlev);
  CK(cudaMemcpy3DAsync(&p, st));
  return 0;
}
static int put(mom6adv_plan *plan, cudaStream_t st, double *dev, const double *src, int ilo, int jlo, int nlev) {
  if (is_device_ptr(src)) return upload(plan, dev, src, ilo, jlo, nlev);      // repack kernel on the main stream
  return xfer3d(plan, st, dev, src, ilo, jlo, nlev, true);
}
static int get(mom6adv_plan *plan, cudaStream_t st, double *dst, const double *dev, int ilo, int jlo, int nlev) {
  if (is_device_ptr(dst)) return download(plan, dst, dev, ilo, jlo, nlev);
  return xfer3d(plan, st, const_cast<double *>(dev), dst, ilo, jlo, nlev, false);
}


// ---------------------------------------------------------------------------
// MIRROR mode: one call == one reference advect_tracer call on host arrays.
// Tracers are independent of one another, so the call is pipelined in chunks:
// H2D of chunk c+1 (stream s_in), the whole algorithm on chunk c (main stream)
// and D2H of chunk c-1 (stream s_out) overlap, and both PCIe directions stay
// busy.  The tracer-independent recursion (uhr, vhr, hprev, flags) is simply
// repeated per chunk from the resident h_end/uhtr/vhtr: a few ms against the
// ~75 ms a chunk spends on the bus.
// ---------------------------------------------------------------------------
static int advect_mirror(mom6adv_plan *plan, const double *h_end, const double *uhtr, const double *vhtr,
                         double dt, int x_first, double *vol_prev, int max_iter_in, int update_vol_prev,
                         double *uhr_out, double *vhr_out, int *n_iter);
extern "C" int mom6adv_advect(mom6adv_plan *plan, const double *h_end, const double *uhtr, const double *vhtr,
                              double dt, int x_first, double *vol_prev, int max_iter_in, int update_vol_prev,
                              double *uhr_out, double *vhr_out, int *n_iter) {
  const int rc = advect_mirror(plan, h_end, uhtr, vhtr, dt, x_first, vol_prev, max_iter_in, update_vol_prev, uhr_out,
                               vhr_out, n_iter);
  if (rc != MOM6ADV_OK && plan) {
    // a failed call must not leave copies in flight against the caller's host arrays: drain the three streams
    // (errors of the drain itself are not reported over the one that brought us here)
    if (plan->s_in) cudaStreamSynchronize(plan->s_in);
    if (plan->s_out) cudaStreamSynchronize(plan->s_out);
    if (plan->stream) cudaStreamSynchronize(plan->stream);
    cudaGetLastError();
  }
  return rc;
}
static int advect_mirror(mom6adv_plan *plan, const double *h_end, const double *uhtr, const double *vhtr,
                         double dt, int x_first, double *vol_prev, int max_iter_in, int update_vol_prev,
                         double *uhr_out, double *vhr_out, int *n_iter) {
  if (n_iter) *n_iter = 0;
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "advect: tracer_advect_init must be called before advect_tracer");
  if (plan->ntr == 0) return MOM6ADV_OK;   // :124
  if (!h_end || !uhtr || !vhtr) return fail(plan, MOM6ADV_ERR_INVALID, "advect: NULL input array");
  bool any_dev = false, any_diag = false;
  for (int m = 0; m < plan->ntr; m++) {
    if (!plan->tr[m].bound || !plan->tr[m].host_t)
      return fail(plan, MOM6ADV_ERR_NOT_INIT, "advect: register_tracer must be called before advect_tracer (tracer not bound)");
    any_dev = any_dev || is_device_ptr(plan->tr[m].host_t);
    for (int d = 0; d < 5; d++) any_diag = any_diag || plan->tr[m].diag[d];
  }
  CK(cudaSetDevice(plan->device));
  const Geom &g = plan->g;
  cudaStream_t s = plan->stream;
  if (!plan->s_in) {
    CK(cudaStreamCreateWithFlags(&plan->s_in, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&plan->s_out, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&plan->ev_huv, cudaEventDisableTiming));
  }
  // device-pointer registries and diagnostic runs are not pipelined (their copies are kernels on the main stream)
  const int chunk = (any_dev || any_diag || plan->opt_chunk <= 0) ? plan->ntr : plan->opt_chunk;
  const int nchunk = (plan->ntr + chunk - 1) / chunk;
  while ((int)plan->ev_chunk.size() < 4 * nchunk) { cudaEvent_t e; CK(cudaEventCreate(&e)); plan->ev_chunk.push_back(e); }   // per chunk: uploaded, compute start, compute end, downloaded
  int rc;
  cudaEvent_t tr_ev[6] = {};
  if (plan->opt_trace) {
    for (auto &e : tr_ev) CK(cudaEventCreate(&e));
    CK(cudaEventRecord(tr_ev[0], s)); CK(cudaEventRecord(tr_ev[2], plan->s_in));
  }
  // h_end, uhtr, vhtr(, vol_prev) go first on the upload stream, so that the copy engine serves them before
  // the first tracer chunk (nothing is being downloaded yet: the strided copy is harmless here)
  {
    cudaStream_t su = any_dev ? s : plan->s_in;
    if (vol_prev && !plan->vol_prev) { rc = dev_alloc(plan, &plan->vol_prev, field_elems(plan)); if (rc) return rc; CK(cudaStreamSynchronize(s)); }
    if ((rc = put(plan, su, plan->h_end, h_end, g.isd, g.jsd, g.nk))) return rc;
    if ((rc = put(plan, su, plan->uhtr, uhtr, g.IsdB, g.jsd, g.nk))) return rc;
    if ((rc = put(plan, su, plan->vhtr, vhtr, g.isd, g.JsdB, g.nk))) return rc;
    if (vol_prev && (rc = put(plan, su, plan->vol_prev, vol_prev, g.isd, g.jsd, g.nk))) return rc;
    if (!any_dev) { CK(cudaEventRecord(plan->ev_huv, plan->s_in)); CK(cudaStreamWaitEvent(s, plan->ev_huv, 0)); }
  }
  // Uploads run `ahead` chunks in front of the compute, as CONTIGUOUS copies into a ring of staging slots
  // that a repack kernel then spreads into the padded layout.  (Measured on this box, scripts/pcie_probe2.py:
  // a row-strided H2D copy slows a concurrent D2H copy to 35 GB/s, a contiguous one leaves it at 50 GB/s;
  // strided D2H costs nothing, so results go straight from the padded layout to the host array.)
  const int ahead = std::max(1, plan->opt_ahead);
  const int nslot = ahead + 1;
  const long long telems = (long long)(g.ied - g.isd + 1) * (g.jed - g.jsd + 1) * g.nk;
  if (!any_dev && plan->stage_in_elems < telems * chunk * nslot) {
    if (plan->stage_in) { CK(cudaStreamSynchronize(s)); cudaFree(plan->stage_in); plan->stage_in = nullptr; plan->dev_bytes -= 8 * plan->stage_in_elems; }
    plan->stage_in_elems = telems * chunk * nslot;
    if ((rc = dev_alloc(plan, &plan->stage_in, plan->stage_in_elems))) return rc;
    CK(cudaStreamSynchronize(s));     // dev_alloc clears on the main stream: must not race the first upload
  }
  auto t_host0 = std::chrono::steady_clock::now();
  double host_enq_ms = 0.0;
  auto enqueue_in = [&](int c) -> int {
    if (c >= nchunk) return 0;
    auto t0 = std::chrono::steady_clock::now();
    const int m0 = c * chunk, m1 = std::min(plan->ntr, m0 + chunk);
    if (any_dev) {
      for (int m = m0; m < m1; m++) {
        int r = put(plan, s, plan->tr[m].buf[plan->tr[m].cur], plan->tr[m].host_t, g.isd, g.jsd, g.nk);
        if (r) return r;
      }
      CK(cudaEventRecord(plan->ev_chunk[4 * c], s));
    } else {
      double *slot = plan->stage_in + (long long)(c % nslot) * chunk * telems;
      if (c >= nslot) CK(cudaStreamWaitEvent(plan->s_in, plan->ev_chunk[4 * (c - nslot) + 1], 0));   // slot repacked
      for (int m = m0; m < m1; m++)
        CK(cudaMemcpyAsync(slot + (long long)(m - m0) * telems, plan->tr[m].host_t, sizeof(double) * telems,
                           cudaMemcpyHostToDevice, plan->s_in));
      CK(cudaEventRecord(plan->ev_chunk[4 * c], plan->s_in));
      if (plan->opt_trace && c == nchunk - 1) CK(cudaEventRecord(tr_ev[3], plan->s_in));
    }
    host_enq_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return 0;
  };
  for (int c = 0; c < std::min(ahead, nchunk); c++) if ((rc = enqueue_in(c))) return rc;
  const bool upd = vol_prev && update_vol_prev > 0;
  double ms_total = 0.0;
  for (int c = 0; c < nchunk; c++) {
    const int m0 = c * chunk, m1 = std::min(plan->ntr, m0 + chunk);
    if ((rc = enqueue_in(c + ahead))) return rc;
    CK(cudaStreamWaitEvent(s, plan->ev_chunk[4 * c], 0));
    if (!any_dev) {
      const double *slot = plan->stage_in + (long long)(c % nslot) * chunk * telems;
      for (int m = m0; m < m1; m++)
        launch_repack_in(s, g, plan->tr[m].buf[plan->tr[m].cur], slot + (long long)(m - m0) * telems, g.isd, g.jsd, g.nk);
      plan->launches += m1 - m0;
    }
    CK(cudaEventRecord(plan->ev_chunk[4 * c + 1], s));        // the slot is free again; compute starts
    rc = run_advect(plan, dt, x_first, max_iter_in, vol_prev != nullptr, upd && c == nchunk - 1, n_iter, m0, m1);
    if (rc) return rc;
    ms_total += plan->last_ms;
    cudaStream_t so = any_dev ? s : plan->s_out;
    CK(cudaEventRecord(plan->ev_chunk[4 * c + 2], s));
    CK(cudaStreamWaitEvent(so, plan->ev_chunk[4 * c + 2], 0));
    if (plan->opt_trace && c == 0) CK(cudaEventRecord(tr_ev[4], so));
    for (int m = m0; m < m1; m++) {
      if ((rc = get(plan, so, plan->tr[m].host_t, plan->tr[m].buf[plan->tr[m].cur], g.isd, g.jsd, g.nk))) return rc;
      for (int d = 0; d < 5; d++) if (plan->tr[m].diag[d] && plan->tr[m].host_diag[d]) {
        int ilo, jlo, nlev;
        diag_info(g, d, &ilo, &jlo, &nlev);
        if ((rc = get(plan, so, plan->tr[m].host_diag[d], plan->tr[m].diag[d], ilo, jlo, nlev))) return rc;
      }
    }
    if (plan->opt_trace) CK(cudaEventRecord(plan->ev_chunk[4 * c + 3], so));
  }
  plan->last_ms = ms_total;
  if (uhr_out && (rc = get(plan, s, uhr_out, plan->uhr, g.IsdB, g.jsd, g.nk))) return rc;   // :316
  if (vhr_out && (rc = get(plan, s, vhr_out, plan->vhr, g.isd, g.JsdB, g.nk))) return rc;   // :317
  if (upd && (rc = get(plan, s, vol_prev, plan->vol_prev, g.isd, g.jsd, g.nk))) return rc;  // :319
  if (plan->opt_trace) { CK(cudaEventRecord(tr_ev[5], any_dev ? s : plan->s_out)); CK(cudaEventRecord(tr_ev[1], s)); }
  CK(cudaStreamSynchronize(plan->s_out));
  CK(cudaStreamSynchronize(s));
  if (plan->opt_trace) {
    float a = 0, b = 0, c2 = 0, d = 0, e = 0;
    cudaEventElapsedTime(&a, tr_ev[0], tr_ev[1]); cudaEventElapsedTime(&b, tr_ev[2], tr_ev[3]);
    cudaEventElapsedTime(&c2, tr_ev[4], tr_ev[5]); cudaEventElapsedTime(&d, tr_ev[0], tr_ev[4]); cudaEventElapsedTime(&e, tr_ev[0], tr_ev[5]);
    fprintf(stderr, "[mom6adv trace] chunks %d x %d tracers: main %.1f ms, H2D stream %.1f ms, D2H stream %.1f ms (starts at %.1f, ends at %.1f), compute sum %.1f ms; host: %.1f ms in H2D enqueues, %.1f ms total\n",
            nchunk, chunk, a, b, c2, d, e, ms_total, host_enq_ms,
            std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_host0).count());
    if (plan->opt_trace > 1)
      for (int c = 0; c < nchunk; c++) {
        float t[4];
        for (int q = 0; q < 4; q++) cudaEventElapsedTime(&t[q], tr_ev[0], plan->ev_chunk[4 * c + q]);
        fprintf(stderr, "[mom6adv trace]   chunk %2d: uploaded %7.1f  compute %7.1f .. %7.1f  downloaded %7.1f ms\n", c, t[0], t[1], t[2], t[3]);
      }
    for (auto &ev : tr_ev) cudaEventDestroy(ev);
  }
  return MOM6ADV_OK;
}

extern "C" int mom6adv_sync(mom6adv_plan *plan) {
  if (!plan) return fail(nullptr, MOM6ADV_ERR_NOT_INIT, "sync: plan not initialised");
  CK(cudaSetDevice(plan->device));
  CK(cudaStreamSynchronize(plan->stream));
  return MOM6ADV_OK;
}

extern "C" double mom6adv_last_advect_ms(const mom6adv_plan *plan) { return plan ? plan->last_ms : 0.0; }
extern "C" long long mom6adv_kernel_launches(const mom6adv_plan *plan) { return plan ? plan->launches : 0; }
extern "C" long long mom6adv_device_bytes(const mom6adv_plan *plan) { return plan ? plan->dev_bytes : 0; }
extern "C" long long mom6adv_halo_bytes(const mom6adv_plan *plan) { return plan ? plan->halo_bytes_sent : 0; }

extern "C" void *mom6adv_stream(const mom6adv_plan *plan) { return plan ? (void *)plan->stream : nullptr; }

extern "C" int mom6adv_kernel_class_time(const mom6adv_plan *plan, int kclass, double *ms_total, long long *n) {
  if (!plan || kclass < 0 || kclass >= MOM6ADV_KC_COUNT) return MOM6ADV_ERR_INVALID;
  if (ms_total) *ms_total = plan->prof_ms[kclass];
  if (n) *n = plan->prof_n[kclass];
  return MOM6ADV_OK;
}

extern "C" int mom6adv_set_option(mom6adv_plan *plan, const char *key, int value) {
  if (!plan || !key) return fail(plan, MOM6ADV_ERR_INVALID, "set_option: NULL argument");
  if (!strcmp(key, "fused")) { plan->opt_fused = value; return MOM6ADV_OK; }
  if (!strcmp(key, "fast")) { plan->opt_fast = value; return MOM6ADV_OK; }
  if (!strcmp(key, "dirty")) { plan->opt_dirty = value; return MOM6ADV_OK; }
  if (!strcmp(key, "pull2")) { plan->opt_pull2 = value; return MOM6ADV_OK; }
  if (!strcmp(key, "sparse")) { plan->opt_sparse = value; return MOM6ADV_OK; }
  if (!strcmp(key, "obc_multi")) { plan->opt_obc_multi = value; return MOM6ADV_OK; }
  if (!strcmp(key, "hordiff_df")) { plan->opt_hordiff_df = value; return MOM6ADV_OK; }
  if (!strcmp(key, "mirror_chunk")) { plan->opt_chunk = value; return MOM6ADV_OK; }
  if (!strcmp(key, "mirror_ahead")) { plan->opt_ahead = value; return MOM6ADV_OK; }
  if (!strcmp(key, "pull")) { if (plan->ipc_ntr >= 0 && value != plan->opt_pull) return fail(plan, MOM6ADV_ERR_INVALID, "set_option: pull must be chosen before the first advect call"); plan->opt_pull = value; return MOM6ADV_OK; }
  if (!strcmp(key, "sparse_cap")) { plan->opt_sp_cap = value; return MOM6ADV_OK; }
  if (!strcmp(key, "overlap")) { plan->opt_overlap = value; return MOM6ADV_OK; }
  if (!strcmp(key, "trace")) { plan->opt_trace = value; return MOM6ADV_OK; }
  if (!strcmp(key, "profile")) {
    plan->opt_profile = value;
    for (int c = 0; c < MOM6ADV_KC_COUNT; c++) { plan->prof_ms[c] = 0.0; plan->prof_n[c] = 0; }
    return MOM6ADV_OK;
  }
  return fail(plan, MOM6ADV_ERR_INVALID, std::string("set_option: unknown key ") + key);
}
