"""GPU probe: host facts + a quick resident-mode timing at a mid-size shape."""
import os, sys, time, subprocess
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
print(subprocess.run("nproc; free -g | head -2; nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv", shell=True, capture_output=True, text=True).stdout)
from mom6_b200 import synth, capi
from mom6_b200.tracer_advect import tracer_advect_init
ni, nj, nk, ntr = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (360, 210, 75, 22))]
t = time.time()
c = synth.make_case(ni, nj, nk, ntr, cfl=0.3, device="cuda")
torch.cuda.synchronize(); print("gen %.2fs" % (time.time() - t))
CS = tracer_advect_init(None, c.G, None, {"DT": 900.0, "TRACER_ADVECTION_SCHEME": "PPM"}, None, None)
p = capi.Plan(c.G, c.GV, CS)
p.set_ntr(ntr)
for m in range(ntr):
    p.upload_tracer(m, c.tracer(m))
p.upload_field(capi.F_H_END, c.h_end); p.upload_field(capi.F_UHTR, c.uhtr); p.upload_field(capi.F_VHTR, c.vhtr)
print("device bytes %.2f GB" % (p.device_bytes / 1e9))
for it in range(5):
    n = p.advect_resident(c.dt)
    ms = p.last_advect_ms
    N = ni * nj * nk
    print("call %d: iters %d  %.3f ms  %.3e cell-tracer/s  alg %.1f GB/s" % (it, n, ms, N * ntr / ms * 1e3, (16 * ntr + 24) * N / ms / 1e6))
