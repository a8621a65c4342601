#!/usr/bin/env python
"""Golden .npz (tests/golden/make_golden.py) -> binary case file for advect_driver.  Test infrastructure.

Layout (little endian): int32 hdr[28] = magic, isc, iec, jsc, jec, isd, ied, jsd, jed, IsdB, IedB, JsdB, JedB, nk,
first_direction, ntr, scheme (0 PLM, 1 PPM, 2 PPM:H3), huynh_stencil_bug, x_first (-1 absent), max_iter_in (-1 absent),
update_vol_prev (-1 absent), has_vol_prev, ncalls, cyclic_x, cyclic_y, tripolar_n, diags, 0;
float64: dt, dt_baroclinic, Angstrom_H, H_subroundoff; areaT, IareaT, mask2dT, mask2dCu, mask2dCv; h_end, uhtr, vhtr;
[vol_prev]; per tracer: conc_underflow, t.  Arrays are C-order [k, j, i] == Fortran (i, j, k)."""
import importlib.util
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
_spec = importlib.util.spec_from_file_location("make_golden", os.path.join(ROOT, "tests", "golden", "make_golden.py"))
mg = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(mg)
MAGIC = 1297041718


def main(npz, out):
    name = os.path.splitext(os.path.basename(npz))[0]
    spec = mg.CASES[name]
    ni, nj, nk, ntr, scheme, cfl, x_first, ncalls, diags, offline, kw = spec
    c, _ = mg.build_case(spec)
    G, D = c.G, c.G.Domain
    z = np.load(npz)
    hdr = np.array([MAGIC, G.isc, G.iec, G.jsc, G.jec, G.isd, G.ied, G.jsd, G.jed, G.IsdB, G.IedB, G.JsdB, G.JedB, nk,
                    G.first_direction, ntr, {"PLM": 0, "PPM": 1, "PPM:H3": 2}[scheme], 0,
                    -1 if x_first is None else int(x_first), 2 if offline else -1, 1 if offline else -1,
                    1 if offline else 0, ncalls, int(D.cyclic_x), int(D.cyclic_y), int(D.tripolar_n), int(diags), 0],
                   dtype="<i4")
    with open(out, "wb") as f:
        f.write(hdr.tobytes())
        f.write(np.array([z["dt"], z["dt_baroclinic"], z["Angstrom_H"], z["H_subroundoff"]], dtype="<f8").tobytes())
        for k in ("G_areaT", "G_IareaT", "G_mask2dT", "G_mask2dCu", "G_mask2dCv", "h_end", "uhtr", "vhtr"):
            f.write(np.ascontiguousarray(z[k], dtype="<f8").tobytes())
        if offline:
            f.write(np.ascontiguousarray(z["vol_prev"], dtype="<f8").tobytes())
        for m in range(ntr):
            f.write(np.array([z["conc_underflow"][m]], dtype="<f8").tobytes())
            f.write(np.ascontiguousarray(z["t_in_%02d" % m], dtype="<f8").tobytes())


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
