"""mom6adv_zint100_noalloc (no counterpart in the reference): the 100 m
column integral on caller-owned device arrays with the library's fixed file-scope scratch and no allocation.

Checked through the C ABI against the existing entry point mom6adv_column_diag(kind = ZINT_100M), bitwise (the
tolerance every test of this repository uses for this path), on two calls in a row that differ in shape and in the
layer bound khi -- so a value the first call left in the shared scratch would show in the second.  Collected last
(file name), independent of the other tests: it builds its own plan and arrays."""
import numpy as np
import pytest

from mom6_b200 import synth
from mom6_b200.tracer_advect import tracer_advect_init

pytestmark = pytest.mark.gpu


def _existing(c, tracer, h_diag, H_to_m):
    """mom6adv_column_diag(ZINT_100M) for one resident tracer: (jsd:jed, isd:ied) array, zero outside the compute domain."""
    from mom6_b200 import capi
    CS = tracer_advect_init(None, c.G, None, {"DT": c.dt_baroclinic, "TRACER_ADVECTION_SCHEME": "PPM"}, None, None)
    p = capi.Plan(c.G, c.GV, CS)
    p.set_ntr(1)
    p.upload_tracer(0, tracer)
    p.bind_tracer(0, None, 0.0)
    out = np.full(c.G.shape_h(), 7.0)
    p.column_diag(0, capi.CD_ZINT_100M, h_diag, H_to_m, out)
    p.destroy()
    return out


def _case(ni, nj, nk, seed, H_to_m):
    c = synth.make_case(ni, nj, nk, 2, cfl=0.4, seed=synth.SEED0 + seed)
    rng = np.random.default_rng(seed)
    h = c.h_end.numpy()
    h_diag = np.ascontiguousarray(h * (1.0 + 0.05 * rng.standard_normal(h.shape)))
    t = np.ascontiguousarray(synth.bind_registry(c).Tr[1].t)
    return c, h_diag, t, H_to_m


def test_no_alloc_entry_equals_column_diag_twice_in_a_row():
    import torch
    from mom6_b200 import capi
    assert torch.cuda.is_available()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    # two different shapes; the second has the SHALLOWER layer bound (large H_to_m: 100 m is reached in the first layers)
    cases = [_case(61, 43, 12, 3, 0.5), _case(37, 29, 9, 5, 40.0)]
    khis = []
    for n, (c, h_diag, t, H_to_m) in enumerate(cases):
        G, nk = c.G, c.GV.ke
        want = G.interior(_existing(c, t, h_diag, H_to_m))
        # the caller's device arrays: the computational domain, (nk, nj, ni) C order == (ni, nj, nk) Fortran order
        hd = torch.from_numpy(np.ascontiguousarray(G.interior(h_diag))).to(dev)
        td = torch.from_numpy(np.ascontiguousarray(G.interior(t))).to(dev)
        out = torch.full((G.nj, G.ni), 7.0, dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        capi.zint100_noalloc(hd, td, G.ni, G.nj, nk, H_to_m, out, stream=0)
        torch.cuda.synchronize()
        free1 = torch.cuda.mem_get_info()[0]
        if n > 0:      # (the very first launch of a kernel may load its module: judge the allocation-free claim on the second call)
            assert free1 == free0, "device memory changed across a no-alloc call: %d -> %d" % (free0, free1)
        got = out.cpu().numpy()
        assert got.shape == want.shape
        assert np.array_equal(got.view(np.uint64), np.ascontiguousarray(want).view(np.uint64)), \
            "call %d differs from mom6adv_column_diag: max abs %g" % (n, np.abs(got - want).max())
        assert np.abs(got).max() > 0.0
        # layer bound of this case (numpy, for the test's own sanity: the two calls must exercise different bounds)
        z = np.cumsum(G.interior(h_diag) * H_to_m, axis=0)
        ztop = z - G.interior(h_diag) * H_to_m
        khis.append(int(np.max(np.where((ztop < 100.0).any(axis=(1, 2)))[0])) + 1)
    assert khis[1] < khis[0], khis


def test_no_alloc_entry_rejects_sizes_beyond_its_scratch():
    import torch
    from mom6_b200 import capi
    a = torch.zeros(8, dtype=torch.float64, device="cuda")
    with pytest.raises(capi.Mom6AdvError) as e:
        capi.zint100_noalloc(a, a, 4096, 4096, 1, 1.0, a)      # 16.8 M columns > 8 388 608
    assert e.value.status == capi.ERR_UNSUPPORTED and "MOM6ADV_NOALLOC_MAX_COLUMNS" in str(e.value)
    with pytest.raises(capi.Mom6AdvError) as e:
        capi.zint100_noalloc(np.zeros(8), a, 2, 2, 2, 1.0, a)  # a host array is refused, not dereferenced
    assert e.value.status == capi.ERR_INVALID
