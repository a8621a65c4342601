#!/usr/bin/env python
"""Compare advect_driver's output with the golden .npz, bitwise over the computational domain.  Test infrastructure."""
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


def main(npz, out):
    name = os.path.splitext(os.path.basename(npz))[0]
    spec = mg.CASES[name]
    ntr, diags, offline = spec[3], spec[8], spec[9]
    c, _ = mg.build_case(spec)
    G, nk = c.G, c.GV.ke
    z = np.load(npz)
    raw = np.fromfile(out, dtype="<f8")
    pos = [0]

    def take(shape):
        n = int(np.prod(shape))
        a = raw[pos[0]:pos[0] + n].reshape(shape)
        pos[0] += n
        return a
    bad = 0

    def check(label, got, want, cut):
        nonlocal bad
        if not np.array_equal(np.ascontiguousarray(cut(got)).view(np.uint64), np.ascontiguousarray(cut(want)).view(np.uint64)):
            bad += 1
            print("DIFF %-18s max abs %g" % (label, np.abs(cut(got) - cut(want)).max()))
    for m in range(ntr):
        check("tracer %d" % m, take(G.shape_h(nk)), z["t_out_%02d" % m], G.interior)
    check("uhr_out", take(G.shape_u(nk)), z["uhr_out"], G.interior_u)
    check("vhr_out", take(G.shape_v(nk)), z["vhr_out"], G.interior_v)
    if offline:
        check("vol_prev", take(G.shape_h(nk)), z["vol_prev_out"], G.interior)
    if diags:
        for m in range(ntr):
            check("ad_x %d" % m, take(G.shape_u(nk)), z["ad_x_%02d" % m], G.interior_u)
            check("ad_y %d" % m, take(G.shape_v(nk)), z["ad_y_%02d" % m], G.interior_v)
            check("ad2d_x %d" % m, take(G.shape_u()), z["ad2d_x_%02d" % m], G.interior_u)
            check("ad2d_y %d" % m, take(G.shape_v()), z["ad2d_y_%02d" % m], G.interior_v)
            check("advection_xy %d" % m, take(G.shape_h(nk)), z["advection_xy_%02d" % m], G.interior)
    assert pos[0] == raw.size, "output file has %d extra values" % (raw.size - pos[0])
    print("BITWISE OK" if bad == 0 else "%d fields differ" % bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1], sys.argv[2]))
