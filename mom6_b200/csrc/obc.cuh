// Device-side tables of the open-boundary path (obc.cu) and its launchers (internal).
#pragma once
#include "internal.cuh"
#include "kernels.h"

enum { OBC_DIR_N = 100, OBC_DIR_S = 200, OBC_DIR_E = 300, OBC_DIR_W = 400 };   // MOM_open_boundary.F90:80-83

struct ObcSeg {            // OBC_segment_type members the path reads
  int dir, is_ew, is_ns, specified, has_reg;
  int isd, ied, jsd, jed, IsdB, IedB, JsdB, JedB;      // segment%HI
  int ntseg, tr_off;       // entries [tr_off, tr_off + ntseg) of the ObcTr table
  long long cond_off;      // first byte of this segment in the cond array
};
struct ObcTr {             // OBC_segment_tracer_type
  int m;                   // registry index, 0-based
  double conc;             // OBC_inflow_conc
  const double *tres;      // device copy of the reservoir (segment bounds), or NULL
};

void launch_obc_slope(cudaStream_t s, int dir, const Geom &g, Range rg, int sten, const double *T, const double *mm,
                      double *slope, const int *domore_k);
void launch_obc_override(cudaStream_t s, int dir, const Geom &g, Range rg, const ObcSeg &sg, const ObcTr *trs, int m,
                         double *Ttmp, double *slope, const double *mm, const int *entry, const int *domore_k);
void launch_obc_mass(cudaStream_t s, int dir, const Geom &g, Range rg, const ObcSeg &sg, int blk, const double *res,
                     double *hh2, const double *maskT, unsigned char *cond, const int *entry, const int *domore_k);
void launch_obc_flux_update(cudaStream_t s, int scheme, int dir, const Geom &g, Range rg, const TracerDev &tr, int m,
                            int src, const double *Ttmp, const double *slope, const double *hh, const double *hh2,
                            const double *cfl, const double *hlst, const double *ihnew, const double *mm,
                            const double *IareaT, const int *entry, const int *domore_k, double Idt,
                            const ObcSeg *segs, int nseg, const ObcTr *trs, const unsigned char *cond, double *fx);
void launch_obc_ad2d(cudaStream_t s, int dir, const Geom &g, Range rg, double *a2, const double *fx,
                     const double *ihnew, const int *entry, const int *domore_k, double Idt);
