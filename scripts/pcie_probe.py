"""PCIe probe: pinned H2D, D2H and both at once (what bounds the MIRROR-mode e2e number).
Per-direction rates when both run, with 1 or 2 streams per direction."""
import torch
n = 1 << 30   # 1 GiB per buffer
NB = 8
h_in = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(NB)]
h_out = [torch.empty(n, dtype=torch.uint8).pin_memory() for _ in range(NB)]
d_in = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(NB)]
d_out = [torch.empty(n, dtype=torch.uint8, device="cuda") for _ in range(NB)]
S = [torch.cuda.Stream() for _ in range(6)]
def run(n_up, n_down, ups, downs):
    """n_up/n_down buffers per direction over `ups`/`downs` streams; returns (GB/s up, GB/s down) over each direction's own span."""
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(2)] for _ in range(6)]
    torch.cuda.synchronize()
    t0 = torch.cuda.Event(enable_timing=True); t0.record()
    for st in S: st.wait_event(t0)
    for i in range(max(n_up, n_down)):
        if i < n_up:
            with torch.cuda.stream(S[i % ups]): d_in[i].copy_(h_in[i], non_blocking=True)
        if i < n_down:
            with torch.cuda.stream(S[3 + i % downs]): h_out[i].copy_(d_out[i], non_blocking=True)
    ends = []
    for k, st in enumerate(S):
        e = torch.cuda.Event(enable_timing=True); e.record(st); ends.append(e)
    torch.cuda.synchronize()
    tu = max(t0.elapsed_time(ends[k]) for k in range(3)) if n_up else 0
    td = max(t0.elapsed_time(ends[k]) for k in range(3, 6)) if n_down else 0
    return (n_up * n / tu / 1e6 if n_up else 0, n_down * n / td / 1e6 if n_down else 0)
run(2, 2, 1, 1)
print("alone: H2D %.1f  D2H %.1f GB/s" % (run(NB, 0, 1, 1)[0], run(0, NB, 1, 1)[1]))
for ups, downs in ((1, 1), (1, 2), (2, 2), (1, 3)):
    u, d = run(NB, NB, ups, downs)
    print("both, %d up / %d down streams: H2D %.1f  D2H %.1f GB/s (each over its own span)" % (ups, downs, u, d))
