import sys, re, numpy as np
sys.path.insert(0, '/root/repo')
from mom6_b200 import synth, decomp
from mom6_b200.types import TracerRegistry, register_tracer
from mom6_b200.tracer_advect import tracer_advect_init
from oracle import oracle
def per_pe(ncall, scheme="PPM", layout=(2,1)):
    c, OBC = synth.make_obc_case()
    params = {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": scheme}
    h, u, v = c.h_end.numpy(), c.uhtr.numpy(), c.vhtr.numpy()
    R0 = synth.bind_registry(c)
    ranks = []
    for pj in range(layout[1]):
        for pi in range(layout[0]):
            Gr = decomp.local_grid(c.G, layout, (pi, pj))
            Rr = TracerRegistry()
            for tr in R0.Tr:
                register_tracer(decomp.cut_h(c.G, Gr, tr.t), Rr, underflow_conc=tr.conc_underflow)
            ranks.append(dict(G=Gr, GV=c.GV, Reg=Rr, h_end=decomp.cut_h(c.G, Gr, h), uhtr=decomp.cut_u(c.G, Gr, u),
                              vhtr=decomp.cut_v(c.G, Gr, v), OBC=decomp.cut_obc(OBC, c.G, Gr)))
    CSo = tracer_advect_init(None, c.G, None, params, None, None)
    for _ in range(ncall):
        oracle.advect_tracer_multi(ranks, c.G.Domain, CSo, c.dt)
    return ranks
for fn, ncall, rank in (("profiles/r02/obc_multi/one_call_two_trips_rank0.txt", 1, 0), ("profiles/r02/obc_multi/two_calls_rank0.txt", 2, 0), ("profiles/r02/obc_multi/two_calls_rank1.txt", 2, 1)):
    ranks = per_pe(ncall)
    G = ranks[rank]["G"]
    n = ok = 0
    for line in open('/root/repo/' + fn):
        m = re.match(r"tracer (\d+) k=(\d+) j=(\d+) i=(\d+) .* want (\S+) got (\S+)", line)
        if not m: continue
        t, k, j, i = (int(x) for x in m.groups()[:4]); got = float(m.group(6))
        val = G.interior(ranks[rank]["Reg"].Tr[t].t)[k, j, i]
        n += 1; ok += int(np.float64(got).view(np.uint64) == np.float64(val).view(np.uint64))
    print(fn, "cells where the GPU differed from the 1-PE run:", n, "-> GPU value equals the per-PE reference run bitwise at", ok)
