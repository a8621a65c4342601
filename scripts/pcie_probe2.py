"""PCIe probe 2: contiguous vs row-strided (2-D pitched) DMA, alone and in both directions at once.
Row = 1448 doubles (OM4_025 data-domain row), device pitch 1472 doubles, as the MIRROR call copies them."""
import ctypes as C, torch, glob, os
rt = None
for pat in ("/usr/local/cuda/lib64/libcudart.so*", os.path.join(os.path.dirname(torch.__file__), "lib", "libcudart*.so*"),
            os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*")):
    for f in sorted(glob.glob(pat)):
        try:
            rt = C.CDLL(f); break
        except OSError:
            pass
    if rt: break
rt.cudaMemcpy2DAsync.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p]
rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
ROW, PITCH, ROWS = 1448 * 8, 1472 * 8, 1088 * 75
NB = 6
h_in = [torch.empty(ROW * ROWS, dtype=torch.uint8).pin_memory() for _ in range(NB)]
h_out = [torch.empty(ROW * ROWS, dtype=torch.uint8).pin_memory() for _ in range(NB)]
d_in = [torch.empty(PITCH * ROWS + 256, dtype=torch.uint8, device="cuda") for _ in range(NB)]
d_out = [torch.empty(PITCH * ROWS + 256, dtype=torch.uint8, device="cuda") for _ in range(NB)]
s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()
H2D, D2H = 1, 2
def up(i, strided, off):
    if strided:
        rc = rt.cudaMemcpy2DAsync(d_in[i].data_ptr() + off, PITCH, h_in[i].data_ptr(), ROW, ROW, ROWS, H2D, s_up.cuda_stream)
    else:
        rc = rt.cudaMemcpyAsync(d_in[i].data_ptr(), h_in[i].data_ptr(), ROW * ROWS, H2D, s_up.cuda_stream)
    assert rc == 0, rc
def dn(i, strided, off):
    if strided:
        rc = rt.cudaMemcpy2DAsync(h_out[i].data_ptr(), ROW, d_out[i].data_ptr() + off, PITCH, ROW, ROWS, D2H, s_dn.cuda_stream)
    else:
        rc = rt.cudaMemcpyAsync(h_out[i].data_ptr(), d_out[i].data_ptr(), ROW * ROWS, D2H, s_dn.cuda_stream)
    assert rc == 0, rc
def run(do_up, do_dn, su, sd, off=96):
    torch.cuda.synchronize()
    t0 = torch.cuda.Event(enable_timing=True); t0.record()
    s_up.wait_event(t0); s_dn.wait_event(t0)
    for i in range(NB):
        if do_up: up(i, su, off)
        if do_dn: dn(i, sd, off)
    eu, ed = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    eu.record(s_up); ed.record(s_dn)
    torch.cuda.synchronize()
    gb = NB * ROW * ROWS / 1e6
    return (gb / t0.elapsed_time(eu) if do_up else 0.0, gb / t0.elapsed_time(ed) if do_dn else 0.0)
run(1, 1, 0, 0)
for su in (0, 1):
    print("alone  H2D %-10s %.1f GB/s   D2H %-10s %.1f GB/s" % ("strided" if su else "contiguous", run(1, 0, su, su)[0],
                                                              "strided" if su else "contiguous", run(0, 1, su, su)[1]))
for su in (0, 1):
    for sd in (0, 1):
        u, d = run(1, 1, su, sd)
        print("both   H2D %-10s %.1f GB/s   D2H %-10s %.1f GB/s" % ("strided" if su else "contiguous", u, "strided" if sd else "contiguous", d))
u, d = run(1, 1, 1, 1, off=0)
print("both strided, device rows 128B-aligned: H2D %.1f  D2H %.1f GB/s" % (u, d))
