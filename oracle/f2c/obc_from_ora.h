// TEST INFRASTRUCTURE ONLY -- ora_obc (oracle/advect_ref.h) -> the ocean_OBC_type stand-in of f2c_runtime.h, shared by
// the wrappers around the translated reference (ref_wrap.cpp, offline_wrap.cpp).
#pragma once
#include <vector>
#include "f2c_runtime.h"
#include "../advect_ref.h"

struct ObcHolder { ocean_obc_type obc; std::vector<segment_tracer_registry_type> segregs; };

inline ocean_obc_type *obc_from_ora(const ora_obc *OBC, ObcHolder &H, int nk) {
  if (!OBC) return nullptr;
  ocean_obc_type &obc = H.obc;
  std::vector<segment_tracer_registry_type> &segregs = H.segregs;
  const Bnd kk = {1, nk};
  obc.obc_pe = OBC->OBC_pe != 0; obc.number_of_segments = OBC->number_of_segments;
  obc.open_u_bcs_exist_globally = OBC->open_u_BCs_exist_globally != 0;
  obc.open_v_bcs_exist_globally = OBC->open_v_BCs_exist_globally != 0;
  obc.specified_u_bcs_exist_globally = OBC->specified_u_BCs_exist_globally != 0;
  obc.specified_v_bcs_exist_globally = OBC->specified_v_BCs_exist_globally != 0;
  const int ns = OBC->number_of_segments;
  obc.segment = Arr<obc_segment_type, 1>::automatic({{1, ns > 0 ? ns : 1}});
  segregs.resize(ns > 0 ? ns : 1);
  for (int n = 1; n <= ns; n++) {
    const ora_obc_segment &s = OBC->segment[n - 1];
    obc_segment_type &d = obc.segment(n);
    d.is_e_or_w = s.is_E_or_W != 0; d.is_n_or_s = s.is_N_or_S != 0; d.specified = s.specified != 0; d.direction = s.direction;
    d.hi.isd = s.isd; d.hi.ied = s.ied; d.hi.jsd = s.jsd; d.hi.jed = s.jed;
    d.hi.isdb = s.IsdB; d.hi.iedb = s.IedB; d.hi.jsdb = s.JsdB; d.hi.jedb = s.JedB;
    if (s.ntseg >= 0) {
      segment_tracer_registry_type &r = segregs[n - 1];
      r.ntseg = s.ntseg;
      r.tr = Arr<obc_segment_tracer_type, 1>::automatic({{1, s.ntseg > 0 ? s.ntseg : 1}});
      for (int m = 1; m <= s.ntseg; m++) {
        obc_segment_tracer_type &t = r.tr(m);
        t.ntr_index = s.Tr[m - 1].ntr_index; t.obc_inflow_conc = s.Tr[m - 1].OBC_inflow_conc;
        // tres has the segment's own bounds: E/W (IsdB:IedB, jsd:jed, nk), N/S (isd:ied, JsdB:JedB, nk)
        if (s.is_E_or_W) t.tres = Arr<double, 3>::view(const_cast<double *>(s.Tr[m - 1].tres), {{s.IsdB, s.IedB}, {s.jsd, s.jed}, kk});
        else t.tres = Arr<double, 3>::view(const_cast<double *>(s.Tr[m - 1].tres), {{s.isd, s.ied}, {s.JsdB, s.JedB}, kk});
      }
      d.tr_reg = &r;
    }
  }
  return &obc;
}
