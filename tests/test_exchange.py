"""Host-side logic of the multi-GPU group pass (no GPU needed).

The exchange plan the CUDA library builds (mom6adv_exchange_list == what
mom6adv_comm_init uploads) is applied here with numpy -- in one process for all
ranks, and across two real ranks with torch.distributed/gloo -- and compared
with the oracle's emulated do_group_pass (oracle/advect_ref.cpp ora_group_pass),
which the layout-invariance tests of test_oracle.py pin to the 1-PE answer."""
import ctypes as C
import os
import socket

import numpy as np
import pytest

from mom6_b200 import synth, decomp, capi, multi
from mom6_b200.types import MOMDomain, OceanGrid, VerticalGrid
from oracle import oracle

NK = 2


def _global_grid(ni, nj, cyclic_x, cyclic_y, tripolar, symmetric, halo):
    dom = MOMDomain(ni, nj, cyclic_x=cyclic_x, cyclic_y=cyclic_y, tripolar_n=tripolar, symmetric=symmetric)
    return OceanGrid(ni, nj, halo=halo, symmetric=symmetric, domain=dom)


def _rank_fields(G, seed):
    """Fortran-layout arrays of one rank: (uhr, vhr, hprev, t0, t1) with rank- and position-unique values."""
    rng = np.random.default_rng(seed)
    return [rng.standard_normal(G.shape_u(NK)), rng.standard_normal(G.shape_v(NK)),
            rng.standard_normal(G.shape_h(NK)), rng.standard_normal(G.shape_h(NK)), rng.standard_normal(G.shape_h(NK))]


def _oracle_pass(grids, GV, dom, fields):
    """ora_group_pass on copies of the per-rank fields."""
    n = len(grids)
    out = [[a.copy() for a in f] for f in fields]
    lib = oracle.lib()
    lib.ora_group_pass.restype = None
    dp = C.POINTER(C.c_double)
    lib.ora_group_pass.argtypes = [C.c_int, C.POINTER(oracle.ora_rank), C.POINTER(oracle.ora_domain),
                                   C.POINTER(dp), C.POINTER(dp), C.POINTER(dp), C.c_int]
    ranks = (oracle.ora_rank * n)()
    keep = []
    for r, G in enumerate(grids):
        ranks[r].G = oracle.make_grid(G, GV)
        ranks[r].ioff, ranks[r].joff = G.ioff, G.joff
        tr = (oracle.ora_tracer * 2)()
        tr[0].t, tr[1].t = oracle._p(out[r][3]), oracle._p(out[r][4])
        keep.append(tr)
        ranks[r].Tr = C.cast(tr, C.POINTER(oracle.ora_tracer))
    arr = lambda k: (dp * n)(*[oracle._p(out[r][k]) for r in range(n)])
    od = oracle.make_domain(dom)
    lib.ora_group_pass(n, ranks, C.byref(od), arr(0), arr(1), arr(2), 2)
    return out


def _to_padded(G, geom, a, pos):
    """Fortran-layout [k, j, i] array -> padded planes [k, R, P] (NaN elsewhere)."""
    P, R, coff, roff = geom
    ilo = G.IsdB if pos == 1 else G.isd
    jlo = G.JsdB if pos == 2 else G.jsd
    p = np.full((a.shape[0], R, P), np.nan)
    p[:, jlo + roff:jlo + roff + a.shape[1], ilo + coff:ilo + coff + a.shape[2]] = a
    return p


def _from_padded(G, geom, p, pos, shape):
    P, R, coff, roff = geom
    ilo = G.IsdB if pos == 1 else G.isd
    jlo = G.JsdB if pos == 2 else G.jsd
    return p[:, jlo + roff:jlo + roff + shape[1], ilo + coff:ilo + coff + shape[2]].copy()


POS_OF_FIELD = [1, 2, 0, 0, 0]     # uhr, vhr, hprev, t0, t1


def _lists(G, GV, ext, nranks):
    """All lists of one rank: self (dst, src, sgn) and per peer (send_src, send_sgn, recv_dst), per stagger."""
    L = {"self": [], "peer": {}}
    geom = None
    for pos in range(3):
        d, geom = capi.exchange_list(G, GV, ext, -1, pos, 0)
        s, _ = capi.exchange_list(G, GV, ext, -1, pos, 1)
        g, _ = capi.exchange_list(G, GV, ext, -1, pos, 2)
        L["self"].append((d, s, g))
    for peer in range(nranks):
        per = []
        for pos in range(3):
            per.append(tuple(capi.exchange_list(G, GV, ext, peer, pos, w)[0] for w in (3, 4, 5, 6, 7)))
        if any(len(x[0]) or len(x[2]) for x in per):
            L["peer"][peer] = per
    return L, geom


def _pack(L, peer, padded):
    """Message to `peer`: [h: hprev,t0,t1][u: uhr][v: vhr], each [field][level][point] (halo.cu / plan.cu order)."""
    parts = []
    for pos, fidx in ((0, (2, 3, 4)), (1, (0,)), (2, (1,))):
        src, sgn = L["peer"][peer][pos][:2]
        for f in fidx:
            flat = padded[f].reshape(padded[f].shape[0], -1)
            parts.append((flat[:, src] * sgn[None, :]).ravel())
    return np.concatenate(parts) if parts else np.zeros(0)


def _unpack(L, peer, padded, msg):
    off = 0
    for pos, fidx in ((0, (2, 3, 4)), (1, (0,)), (2, (1,))):
        dst = L["peer"][peer][pos][2]
        for f in fidx:
            flat = padded[f].reshape(padded[f].shape[0], -1)
            n = flat.shape[0] * len(dst)
            flat[:, dst] = msg[off:off + n].reshape(flat.shape[0], len(dst))
            off += n
    assert off == msg.size


def _pull(L, peer, padded, peer_padded):
    """Peer-pull form (plan.cu group_pass, k_halo_pull): read the owner's padded planes directly."""
    for f, pos in enumerate(POS_OF_FIELD):
        dst, psrc, psgn = L["peer"][peer][pos][2:5]
        mine = padded[f].reshape(padded[f].shape[0], -1)
        theirs = peer_padded[f].reshape(peer_padded[f].shape[0], -1)
        mine[:, dst] = theirs[:, psrc] * psgn[None, :]


def _apply_self(L, padded):
    for f, pos in enumerate(POS_OF_FIELD):
        d, s, g = L["self"][pos]
        flat = padded[f].reshape(padded[f].shape[0], -1)
        flat[:, d] = flat[:, s] * g[None, :]


def _setup(layout, topo, symmetric=True, halo=4, ni=23, nj=19):
    cyclic_x, cyclic_y, tripolar = topo
    Gg = _global_grid(ni, nj, cyclic_x, cyclic_y, tripolar, symmetric, halo)
    GV = VerticalGrid.default(NK)
    nr = layout[0] * layout[1]
    grids = [decomp.local_grid(Gg, layout, multi.rank_to_pe(r, layout)) for r in range(nr)]
    ext = multi.layout_extents(ni, nj, layout)
    fields = [_rank_fields(G, 100 + r) for r, G in enumerate(grids)]
    return Gg, GV, grids, ext, fields


TOPOS = [(False, False, False), (True, False, False), (False, True, False), (True, True, False), (True, False, True)]


@pytest.mark.parametrize("layout", [(1, 1), (2, 1), (1, 2), (2, 2), (3, 2), (1, 4)])
@pytest.mark.parametrize("topo", TOPOS)
@pytest.mark.parametrize("symmetric", [True, False])
def test_exchange_plan_equals_oracle_group_pass(layout, topo, symmetric):
    if topo[2] and layout[0] > 1 and 23 % layout[0]:
        pytest.skip("fold partners need equal column blocks in this emulation")
    Gg, GV, grids, ext, fields = _setup(layout, topo, symmetric)
    nr = len(grids)
    want = _oracle_pass(grids, GV, Gg.Domain, fields)
    lists, padded = [], []
    for r, G in enumerate(grids):
        L, geom = _lists(G, GV, ext, nr)
        lists.append((L, geom))
        padded.append([_to_padded(G, geom, a, POS_OF_FIELD[f]) for f, a in enumerate(fields[r])])
    msgs = {}
    for r in range(nr):
        for peer in lists[r][0]["peer"]:
            msgs[(r, peer)] = _pack(lists[r][0], peer, padded[r])
    for r in range(nr):
        _apply_self(lists[r][0], padded[r])
        for peer in lists[r][0]["peer"]:
            _unpack(lists[r][0], peer, padded[r], msgs[(peer, r)])
    for r, G in enumerate(grids):
        for f in range(5):
            got = _from_padded(G, lists[r][1], padded[r][f], POS_OF_FIELD[f], fields[r][f].shape)
            assert np.array_equal(got.view(np.uint64), want[r][f].view(np.uint64)), (r, f)
    # the peer-pull form of the same exchange: every rank reads its neighbours' planes as they were before the pass
    before = [[_to_padded(G, lists[r][1], a, POS_OF_FIELD[f]) for f, a in enumerate(fields[r])] for r, G in enumerate(grids)]
    after = [[a.copy() for a in before[r]] for r in range(nr)]
    for r in range(nr):
        _apply_self(lists[r][0], after[r])
        for peer in lists[r][0]["peer"]:
            _pull(lists[r][0], peer, after[r], before[peer])
    for r, G in enumerate(grids):
        for f in range(5):
            got = _from_padded(G, lists[r][1], after[r][f], POS_OF_FIELD[f], fields[r][f].shape)
            assert np.array_equal(got.view(np.uint64), want[r][f].view(np.uint64)), ("pull", r, f)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _gloo_worker(rank, world, port, layout, topo, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        Gg, GV, grids, ext, fields = _setup(layout, topo)
        want = _oracle_pass(grids, GV, Gg.Domain, fields)[rank]
        G = grids[rank]
        L, geom = _lists(G, GV, ext, world)
        padded = [_to_padded(G, geom, a, POS_OF_FIELD[f]) for f, a in enumerate(fields[rank])]
        # the NCCL id bootstrap path (any backend): every rank must end up with the same 128 bytes
        ident = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            ident = torch.arange(128, dtype=torch.uint8)
        dist.broadcast(ident, src=0)
        assert ident[5].item() == 5
        sends = {p: torch.from_numpy(_pack(L, p, padded)) for p in L["peer"]}
        recvs = {p: torch.empty(NK * sum((3 if pos == 0 else 1) * len(L["peer"][p][pos][2]) for pos in range(3)),
                                dtype=torch.float64) for p in L["peer"]}
        reqs = [dist.isend(sends[p], dst=p) for p in sends] + [dist.irecv(recvs[p], src=p) for p in recvs]
        _apply_self(L, padded)
        for r in reqs:
            r.wait()
        for p in recvs:
            _unpack(L, p, padded, recvs[p].numpy())
        ok = all(np.array_equal(_from_padded(G, geom, padded[f], POS_OF_FIELD[f], fields[rank][f].shape).view(np.uint64),
                                want[f].view(np.uint64)) for f in range(5))
        # convergence flags: sum_across_PEs(domore_k, nz) (:305)
        dk = torch.tensor([rank, 0, 1], dtype=torch.int32)
        dist.all_reduce(dk)
        ok = ok and dk.tolist() == [world * (world - 1) // 2, 0, world]
        q.put((rank, ok))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("layout,topo", [((2, 1), (True, False, False)), ((1, 2), (True, False, True)),
                                         ((1, 2), (False, True, False))])
def test_two_rank_exchange_over_gloo(layout, topo):
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, layout, topo, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]


def test_layout_choice_and_extents():
    assert multi.choose_layout(1, 1440, 1080) == (1, 1)
    assert multi.choose_layout(8, 1440, 1080) == (1, 8)
    assert multi.choose_layout(8, 2880, 256)[0] * multi.choose_layout(8, 2880, 256)[1] == 8
    ib, ie, jb, je = multi.layout_extents(10, 7, (3, 2))
    assert ib == [1, 5, 8] and ie == [4, 7, 10] and jb == [1, 5] and je == [4, 7]


# ---- tracer_hordiff's iteration count on two real ranks: max_across_PEs(max_CFL), MOM_tracer_hor_diff.F90:377 --------
def _gloo_num_itts_worker(rank, world, port, layout, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from mom6_b200 import synth, decomp, tracer_hor_diff
        c = synth.make_case(40, 30, 2, 1, cfl=0.4)
        kx, ky = synth.make_khdt(c, 0.05)
        # one strong spot in rank 1's block only: rank 0 alone would ask for fewer iterations
        a = np.asarray(c.G.areaT)
        j, i = c.G.pj(c.G.jec - 2), c.G.pi(c.G.iec - 2)
        kx[j, c.G.pI(c.G.iec - 2)] = 1.3 * a[j, i]
        G = decomp.local_grid(c.G, layout, multi.rank_to_pe(rank, layout))
        n = tracer_hor_diff.num_itts_from_cfl(decomp.cut_u(c.G, G, kx), decomp.cut_v(c.G, G, ky), G)
        q.put((rank, n))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("layout", [(2, 1), (1, 2)])
def test_num_itts_is_the_maximum_over_ranks(layout):
    import torch.multiprocessing as mp
    from mom6_b200 import synth, tracer_hor_diff
    c = synth.make_case(40, 30, 2, 1, cfl=0.4)
    kx, ky = synth.make_khdt(c, 0.05)
    a = np.asarray(c.G.areaT)
    j, i = c.G.pj(c.G.jec - 2), c.G.pi(c.G.iec - 2)
    kx[j, c.G.pI(c.G.iec - 2)] = 1.3 * a[j, i]
    want = tracer_hor_diff.num_itts_from_cfl(kx, ky, c.G)
    assert want >= 3
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_gloo_num_itts_worker, args=(r, 2, port, layout, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
    assert res == [(0, want), (1, want)], (res, want)
