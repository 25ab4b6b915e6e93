"""CPU, world_size 2 over gloo: the N>1 host path — process grid, sub-boxes, the 6-swap schedule and
the staged borders exchange driven by it — with real message passing.  The device kernels are replaced
by numpy selections here; what is under test is the schedule logic of csrc/host.cpp (b200md_procgrid,
b200md_local_box, b200md_comm_schedule) that comm.cu executes on the GPU."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import b200pkg
    B = b200pkg.load().b200md
    scale, n = 3.615, (8, 4, 4)
    boxlo, boxhi = [0.0] * 3, [n[0] * scale, n[1] * scale, n[2] * scale]
    prd = np.array(boxhi) - np.array(boxlo)
    cut = 5.95
    grid = B.procgrid(world, prd)
    loc = B.rank2loc(rank, grid)
    sublo, subhi = B.local_box(boxlo, boxhi, grid, loc)
    x, v, tag = B.lattice_fcc(scale, *n, sublo, subhi, 376847, 63.55)
    cnt = torch.tensor([len(tag)], dtype=torch.int64)
    dist.all_reduce(cnt)
    assert int(cnt) == 4 * n[0] * n[1] * n[2]
    s = B.comm_schedule(boxlo, boxhi, [1, 1, 1], grid, loc, cut)
    # CommBrick::borders driven by the schedule (comm_brick.cpp:771-979)
    X, T = x.copy(), tag.astype(np.int64)
    nlocal = len(T)
    iswap = 0
    for dim in range(3):
        nlast = 0
        for ineed in range(2 * s.maxneed[dim]):
            if ineed % 2 == 0:
                nfirst, nlast = nlast, len(T)
            sel = np.nonzero((X[nfirst:nlast, dim] >= s.slablo[iswap]) & (X[nfirst:nlast, dim] <= s.slabhi[iswap]))[0] + nfirst \
                if s.sendflag[iswap] else np.empty(0, np.int64)
            buf = X[sel].copy()
            for d in range(3):
                buf[:, d] += s.pbc[iswap][d] * prd[d]
            payload = torch.from_numpy(np.concatenate([buf, T[sel, None].astype(np.float64)], axis=1).copy())
            if s.sendproc[iswap] == rank:
                recv = payload
            else:
                nsend = torch.tensor([payload.shape[0]]); nrecv = torch.zeros(1, dtype=torch.int64)
                reqs = [dist.isend(nsend, s.sendproc[iswap]), dist.irecv(nrecv, s.recvproc[iswap])]
                [r.wait() for r in reqs]
                recv = torch.zeros((int(nrecv), 4), dtype=torch.float64)
                reqs = [dist.isend(payload, s.sendproc[iswap]), dist.irecv(recv, s.recvproc[iswap])]
                [r.wait() for r in reqs]
            r = recv.numpy()
            X = np.concatenate([X, r[:, :3]]); T = np.concatenate([T, r[:, 3].astype(np.int64)])
            iswap += 1
    ghosts = X[nlocal:]
    # brute force: every periodic image of every atom inside the extended sub-box, minus owned
    xa, va, ta = B.lattice_fcc(scale, *n, boxlo, [1e30] * 3, 376847, 63.55)
    want = 0
    for sx in (-1, 0, 1):
        for sy in (-1, 0, 1):
            for sz in (-1, 0, 1):
                img = xa + np.array([sx, sy, sz]) * prd
                inside = np.all((img >= np.array(sublo) - cut) & (img <= np.array(subhi) + cut), axis=1)
                own = np.all((img >= np.array(sublo)) & (img < np.array(subhi)), axis=1) & (sx == 0 and sy == 0 and sz == 0)
                want += int(np.sum(inside & ~own))
    out[rank] = (len(ghosts), want, list(grid), int(s.nswap))
    dist.destroy_process_group()


def _xchg(payload, width, sendproc, recvproc, rank):
    """one swap's message: to sendproc, from recvproc (self swaps are a copy)"""
    payload = torch.from_numpy(np.ascontiguousarray(payload, dtype=np.float64).reshape(-1, width))
    if sendproc == rank:
        return payload.numpy().copy()
    nsend = torch.tensor([payload.shape[0]]); nrecv = torch.zeros(1, dtype=torch.int64)
    [r.wait() for r in (dist.isend(nsend, sendproc), dist.irecv(nrecv, recvproc))]
    recv = torch.zeros((int(nrecv), width), dtype=torch.float64)
    [r.wait() for r in (dist.isend(payload, sendproc), dist.irecv(recv, recvproc))]
    return recv.numpy()


def _worker_direct(rank, world, port, out, n=(8, 4, 4)):
    """the direct halo of csrc/comm.cu (B200MD_HALO=direct) restated over gloo: the staged borders carry
    every ghost's owner (rank, index, shift code); forward / reverse traffic then goes owner <-> holder in
    one all-to-all and must reproduce the staged forward_comm bit for bit and the staged reverse_comm sums"""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import b200pkg
    B = b200pkg.load().b200md
    scale = 3.615
    boxlo, boxhi = [0.0] * 3, [n[0] * scale, n[1] * scale, n[2] * scale]
    prd = np.array(boxhi) - np.array(boxlo)
    cut = 5.95
    grid = B.procgrid(world, prd)
    loc = B.rank2loc(rank, grid)
    sublo, subhi = B.local_box(boxlo, boxhi, grid, loc)
    x, v, tag = B.lattice_fcc(scale, *n, sublo, subhi, 376847, 63.55)
    s = B.comm_schedule(boxlo, boxhi, [1, 1, 1], grid, loc, cut)
    nlocal = len(tag)
    X, T = x.copy(), tag.astype(np.int64)
    own_rank = np.full(nlocal, rank, np.int64); own_idx = np.arange(nlocal, dtype=np.int64)
    code = np.zeros((nlocal, 3), np.int64)
    swaps = []
    iswap = 0
    for dim in range(3):
        nlast = 0
        for ineed in range(2 * s.maxneed[dim]):
            if ineed % 2 == 0:
                nfirst, nlast = nlast, len(T)
            sel = np.nonzero((X[nfirst:nlast, dim] >= s.slablo[iswap]) & (X[nfirst:nlast, dim] <= s.slabhi[iswap]))[0] + nfirst \
                if s.sendflag[iswap] else np.empty(0, np.int64)
            pbc = np.array([s.pbc[iswap][d] for d in range(3)], np.int64)
            buf = X[sel].copy()
            for d in range(3):
                buf[:, d] += pbc[d] * prd[d]
            pay = np.concatenate([buf, T[sel, None], own_rank[sel, None], own_idx[sel, None], code[sel] + pbc], axis=1)
            r = _xchg(pay, 9, s.sendproc[iswap], s.recvproc[iswap], rank)
            swaps.append(dict(sendlist=sel, first=len(T), nrecv=len(r), sendproc=s.sendproc[iswap], recvproc=s.recvproc[iswap], pbc=pbc))
            X = np.concatenate([X, r[:, :3]]); T = np.concatenate([T, r[:, 3].astype(np.int64)])
            own_rank = np.concatenate([own_rank, r[:, 4].astype(np.int64)]); own_idx = np.concatenate([own_idx, r[:, 5].astype(np.int64)])
            code = np.concatenate([code, r[:, 6:9].astype(np.int64)])
            iswap += 1
    nall = len(T)
    # the owned atoms move (a deterministic function of the tag), ghosts are stale
    rng_shift = lambda t: 0.05 * np.stack([np.sin(t * 0.37), np.cos(t * 0.11), np.sin(t * 0.23 + 1.0)], axis=1)
    Xn = X.copy(); Xn[:nlocal] += rng_shift(T[:nlocal].astype(np.float64))
    # --- staged forward_comm (comm_brick.cpp:538-591) ---
    Xs = Xn.copy()
    for sw in swaps:
        buf = Xs[sw["sendlist"]].copy()
        for d in range(3):
            buf[:, d] += sw["pbc"][d] * prd[d]
        r = _xchg(buf, 3, sw["sendproc"], sw["recvproc"], rank)
        Xs[sw["first"]:sw["first"] + sw["nrecv"]] = r
    # --- direct: requests grouped by owner rank, lists exchanged once, one all-to-all per step ---
    g = np.arange(nlocal, nall)
    req = {p: g[own_rank[g] == p] for p in range(world)}
    serve = {}
    for p in range(world):
        if p == rank:
            continue
        cnt_out = torch.tensor([len(req[p])]); cnt_in = torch.zeros(1, dtype=torch.int64)
        [r.wait() for r in (dist.isend(cnt_out, p), dist.irecv(cnt_in, p))]
        lst = torch.zeros(int(cnt_in), dtype=torch.int64)
        [r.wait() for r in (dist.isend(torch.from_numpy(own_idx[req[p]].copy()), p), dist.irecv(lst, p))]
        serve[p] = lst.numpy()
    Xd = Xn.copy()
    for p in range(world):
        if p == rank:
            src = Xn[own_idx[req[p]]]
        else:
            recv = torch.zeros((len(req[p]), 3), dtype=torch.float64)
            [r.wait() for r in (dist.isend(torch.from_numpy(Xn[serve[p]].copy()), p), dist.irecv(recv, p))]
            src = recv.numpy()
        q = src.copy()
        for d in range(3):          # one add of prd per crossed image, like the stages (comm.cu shift_n)
            c = code[req[p], d]
            for k in range(1, int(np.abs(c).max(initial=0)) + 1):
                q[:, d] = np.where(c >= k, q[:, d] + prd[d], np.where(-c >= k, q[:, d] - prd[d], q[:, d]))
        Xd[req[p]] = q
    same_forward = bool(np.array_equal(Xd, Xs))
    # --- reverse: every copy holds a value; owners must receive the same sums either way ---
    W = np.zeros(nall); W[nlocal:] = np.cos(T[nlocal:] * 0.7 + rank) + 0.1 * np.arange(nall - nlocal)
    Ws = W.copy()
    for sw in reversed(swaps):
        r = _xchg(Ws[sw["first"]:sw["first"] + sw["nrecv"]], 1, sw["recvproc"], sw["sendproc"], rank)[:, 0]
        np.add.at(Ws, sw["sendlist"], r)
    Wd = W.copy()
    for p in range(world):
        if p == rank:
            np.add.at(Wd, own_idx[req[p]], W[req[p]])
        else:
            recv = torch.zeros(len(serve[p]), dtype=torch.float64)
            [r.wait() for r in (dist.isend(torch.from_numpy(W[req[p]].copy()), p), dist.irecv(recv, p))]
            np.add.at(Wd, serve[p], recv.numpy())
    rev_err = float(np.abs(Wd[:nlocal] - Ws[:nlocal]).max() / np.abs(Ws[:nlocal]).max())
    npeers = sum(1 for p in range(world) if p != rank and len(req[p]))
    out[rank] = (same_forward, rev_err, nall - nlocal, int(sum(len(v) for v in serve.values())),
                 int((code != 0).any(axis=1).sum()), npeers, list(grid))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n,grid,peers", [(2, (8, 4, 4), [2, 1, 1], 1), (4, (4, 4, 4), [1, 2, 2], 3),
                                                (8, (4, 4, 4), [2, 2, 2], 7)])
def test_direct_halo_over_gloo(world, n, grid, peers):
    """2x1x1 (one peer, self swaps in y and z), 1x2x2 and 2x2x2 (the 8-GPU grid: 7 peers, ghosts of
    ghosts owned by third ranks)"""
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_direct, args=(world, 29543 + world, out, n), nprocs=world, join=True)
    assert len(out) == world
    for r in range(world):
        same_forward, rev_err, nghost, nserve, nshifted, npeers, g = out[r]
        assert g == grid and npeers == peers
        assert nghost > 0 and nserve > 0 and nshifted > 0
        assert same_forward, r            # ghost positions bit-identical to the staged forward_comm
        assert rev_err < 1e-12, (r, rev_err)


def test_two_rank_borders_over_gloo():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, 29533, out), nprocs=world, join=True)
    assert len(out) == 2
    for r in range(world):
        got, want, grid, nswap = out[r]
        assert grid == [2, 1, 1] and nswap == 6
        assert got == want, (r, got, want)


def _worker_h26(rank, world, port, out, n=(4, 4, 4)):
    """the direct halo of csrc/halo26.cu restated over gloo.  Its direction tables come from the library's own host
    helper (b200md_halo_directions); the kernels' selections are numpy here.  Checked against the staged borders /
    forward / reverse of CommBrick (restated above) on the same atoms: identical ghost multiset (tag, x, y, z) bit for
    bit, identical refreshed positions, equal reverse sums; and the one-hop migration against brute-force ownership."""
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import b200pkg
    B = b200pkg.load().b200md
    scale = 3.615
    boxlo, boxhi = [0.0] * 3, [n[0] * scale, n[1] * scale, n[2] * scale]
    prd = np.array(boxhi) - np.array(boxlo)
    cut = 5.95
    grid = B.procgrid(world, prd)
    loc = B.rank2loc(rank, grid)
    sublo, subhi = B.local_box(boxlo, boxhi, grid, loc)
    x, v, tag = B.lattice_fcc(scale, *n, sublo, subhi, 376847, 63.55)
    x = x + 0.05 * np.stack([np.sin(tag * 0.37), np.cos(tag * 0.11), np.sin(tag * 0.23 + 1.0)], axis=1)   # off-lattice, inside
    s = B.comm_schedule(boxlo, boxhi, [1, 1, 1], grid, loc, cut)
    nlocal = len(tag)
    T0 = tag.astype(np.int64)

    # ---- staged borders (reference order) ----
    X, T = x.copy(), T0.copy()
    swaps = []
    iswap = 0
    for dim in range(3):
        nlast = 0
        for ineed in range(2 * s.maxneed[dim]):
            if ineed % 2 == 0:
                nfirst, nlast = nlast, len(T)
            sel = np.nonzero((X[nfirst:nlast, dim] >= s.slablo[iswap]) & (X[nfirst:nlast, dim] <= s.slabhi[iswap]))[0] + nfirst \
                if s.sendflag[iswap] else np.empty(0, np.int64)
            pbc = np.array([s.pbc[iswap][d] for d in range(3)], np.int64)
            buf = X[sel].copy()
            for d in range(3):
                buf[:, d] += pbc[d] * prd[d]
            r = _xchg(np.concatenate([buf, T[sel, None]], axis=1), 4, s.sendproc[iswap], s.recvproc[iswap], rank)
            swaps.append(dict(sendlist=sel, first=len(T), nrecv=len(r), sendproc=s.sendproc[iswap], recvproc=s.recvproc[iswap], pbc=pbc))
            X = np.concatenate([X, r[:, :3]]); T = np.concatenate([T, r[:, 3].astype(np.int64)])
            iswap += 1

    # ---- direct: tables, classification, one message per (peer, direction) ----
    peer, shift, shl, slr, usable = B.halo_directions(boxlo, boxhi, [1, 1, 1], grid, loc, cut)
    assert usable
    comp = lambda d: (d % 3 - 1, (d // 3) % 3 - 1, d // 9 - 1)

    def source(d):          # the rank whose copies sent in direction d land here
        l = [(loc[k] - comp(d)[k]) % grid[k] for k in range(3)]
        return (l[0] * grid[1] + l[1]) * grid[2] + l[2]
    peers = sorted({p for p in peer if p >= 0 and p != rank})
    dirs_to = {p: [d for d in range(27) if peer[d] == p] for p in peers}
    dirs_from = {p: [d for d in range(27) if d != 13 and peer[d] >= 0 and source(d) == p] for p in peers}
    dirs_self = [d for d in range(27) if peer[d] == rank]

    def classify(pos):
        L = [pos[:, k] <= shl[k] for k in range(3)]
        R = [pos[:, k] >= slr[k] for k in range(3)]
        lists = {}
        for d in range(27):
            if d == 13 or peer[d] < 0:
                continue
            m = np.ones(len(pos), bool)
            for k in range(3):
                if comp(d)[k] < 0:
                    m &= L[k]
                elif comp(d)[k] > 0:
                    m &= R[k]
            lists[d] = np.nonzero(m)[0]
        return lists

    def shifted(pos, d):
        q = pos.copy()
        for k in range(3):
            q[:, k] += shift[d][k] * prd[k]        # one addition per crossed image, like AtomVec::pack_border
        return q
    lists = classify(x)
    send_idx = {p: np.concatenate([lists[d] for d in dirs_to[p]]) for p in peers}
    send_dir = {p: np.concatenate([np.full(len(lists[d]), d) for d in dirs_to[p]]) for p in peers}

    def pack(pos, p):
        return np.concatenate([shifted(pos[lists[d]], d) for d in dirs_to[p]]) if dirs_to[p] else np.zeros((0, 3))
    # header round (counts per direction), then the payload
    recv_cnt = {}
    for p in peers:
        h_out = torch.tensor([len(lists[d]) for d in dirs_to[p]] + [0] * (32 - len(dirs_to[p])))
        h_in = torch.zeros(32, dtype=torch.int64)
        [r.wait() for r in (dist.isend(h_out, p), dist.irecv(h_in, p))]
        recv_cnt[p] = int(h_in[:len(dirs_from[p])].sum())
    gx, gt = [], []
    for p in peers:
        pay = torch.from_numpy(np.concatenate([pack(x, p), T0[send_idx[p], None].astype(np.float64)], axis=1).copy())
        recv = torch.zeros((recv_cnt[p], 4), dtype=torch.float64)
        [r.wait() for r in (dist.isend(pay, p), dist.irecv(recv, p))]
        gx.append(recv.numpy()[:, :3]); gt.append(recv.numpy()[:, 3].astype(np.int64))
    for d in dirs_self:
        gx.append(shifted(x[lists[d]], d)); gt.append(T0[lists[d]])
    GX, GT = np.concatenate(gx), np.concatenate(gt)
    key = lambda t, p: np.array(sorted(map(tuple, np.concatenate([t[:, None].astype(np.float64), p], axis=1))))
    same_ghosts = len(GT) == len(T) - nlocal and bool(np.array_equal(key(GT, GX), key(T[nlocal:], X[nlocal:])))

    # ---- forward after a move: staged stages vs one direct message per peer ----
    move = lambda t: 0.03 * np.stack([np.cos(t * 0.41), np.sin(t * 0.13), np.cos(t * 0.29 + 2.0)], axis=1)
    xn = x + move(T0.astype(np.float64))
    Xs = np.concatenate([xn, X[nlocal:]])
    for sw in swaps:
        buf = Xs[sw["sendlist"]].copy()
        for d in range(3):
            buf[:, d] += sw["pbc"][d] * prd[d]
        Xs[sw["first"]:sw["first"] + sw["nrecv"]] = _xchg(buf, 3, sw["sendproc"], sw["recvproc"], rank)
    gx = []
    for p in peers:
        recv = torch.zeros((recv_cnt[p], 3), dtype=torch.float64)
        [r.wait() for r in (dist.isend(torch.from_numpy(pack(xn, p).copy()), p), dist.irecv(recv, p))]
        gx.append(recv.numpy())
    for d in dirs_self:
        gx.append(shifted(xn[lists[d]], d))
    same_forward = bool(np.array_equal(key(GT, np.concatenate(gx)), key(T[nlocal:], Xs[nlocal:])))

    # ---- reverse: a value on every ghost copy, summed into the owners ----
    val = lambda t, p: np.cos(t * 0.7) + p[:, 0] * 0.01 + p[:, 1] * 0.02 - p[:, 2] * 0.005     # depends on the copy only
    Ws = np.zeros(len(T)); Ws[nlocal:] = val(T[nlocal:].astype(np.float64), X[nlocal:])
    for sw in reversed(swaps):
        r = _xchg(Ws[sw["first"]:sw["first"] + sw["nrecv"]], 1, sw["recvproc"], sw["sendproc"], rank)[:, 0]
        np.add.at(Ws, sw["sendlist"], r)
    Wd = np.zeros(nlocal)
    gval = val(GT.astype(np.float64), GX)
    off = 0
    for p in peers:
        recv = torch.zeros(len(send_idx[p]), dtype=torch.float64)
        [r.wait() for r in (dist.isend(torch.from_numpy(gval[off:off + recv_cnt[p]].copy()), p), dist.irecv(recv, p))]
        np.add.at(Wd, send_idx[p], recv.numpy())
        off += recv_cnt[p]
    for d in dirs_self:
        np.add.at(Wd, lists[d], gval[off:off + len(lists[d])])
        off += len(lists[d])
    rev_err = float(np.abs(Wd - Ws[:nlocal]).max() / np.abs(Ws[:nlocal]).max())

    # ---- migration: atoms drift up to 1.2 A, are wrapped (Domain::pbc), and move to the neighbour that contains them ----
    xm = x + 1.2 * np.stack([np.sin(T0 * 1.7), np.cos(T0 * 2.3), np.sin(T0 * 0.9 + 0.5)], axis=1)
    xm = np.where(xm < 0, xm + prd, xm); xm = np.where(xm >= prd, np.maximum(xm - prd, 0.0), xm)
    sd = np.zeros((nlocal, 3), np.int64)
    for k in range(3):
        l = [list(loc), list(loc)]
        l[0][k] = (loc[k] - 1) % grid[k]; l[1][k] = (loc[k] + 1) % grid[k]
        (llo, lhi), (rlo, rhi) = B.local_box(boxlo, boxhi, grid, l[0]), B.local_box(boxlo, boxhi, grid, l[1])
        out_k = (xm[:, k] < sublo[k]) | (xm[:, k] >= subhi[k])
        left = out_k & (xm[:, k] >= llo[k]) & (xm[:, k] < lhi[k]) & (grid[k] > 1)
        right = out_k & ~left & (xm[:, k] >= rlo[k]) & (xm[:, k] < rhi[k]) & (grid[k] > 1)
        assert not np.any(out_k & ~left & ~right)
        sd[:, k] = np.where(left, -1, np.where(right, 1, 0))
    dcode = (sd[:, 0] + 1) + 3 * (sd[:, 1] + 1) + 9 * (sd[:, 2] + 1)
    keep = dcode == 13
    arrivals = []
    for p in peers:
        h_out = torch.tensor([int(np.sum(dcode == d)) for d in dirs_to[p]] + [0] * (32 - len(dirs_to[p])))
        h_in = torch.zeros(32, dtype=torch.int64)
        [r.wait() for r in (dist.isend(h_out, p), dist.irecv(h_in, p))]
        selp = np.concatenate([np.nonzero(dcode == d)[0] for d in dirs_to[p]]) if dirs_to[p] else np.empty(0, np.int64)
        pay = torch.from_numpy(np.concatenate([xm[selp], T0[selp, None].astype(np.float64)], axis=1).copy())
        recv = torch.zeros((int(h_in[:len(dirs_from[p])].sum()), 4), dtype=torch.float64)
        [r.wait() for r in (dist.isend(pay, p), dist.irecv(recv, p))]
        arrivals.append(recv.numpy())
    assert all(peer[d] >= 0 and peer[d] != rank for d in set(dcode[~keep].tolist()))
    xo = np.concatenate([xm[keep]] + [a[:, :3] for a in arrivals]); to = np.concatenate([T0[keep]] + [a[:, 3].astype(np.int64) for a in arrivals])
    inside = bool(np.all((xo >= np.array(sublo)) & (xo < np.array(subhi))))
    tot = torch.tensor([len(to)]); dist.all_reduce(tot)
    out[rank] = (same_ghosts, same_forward, rev_err, len(GT), len(peers), inside, int(tot), int((~keep).sum()), list(grid))
    dist.destroy_process_group()


@pytest.mark.parametrize("world,n,grid,peers", [(2, (8, 4, 4), [2, 1, 1], 1), (4, (4, 4, 4), [1, 2, 2], 3),
                                                (8, (4, 4, 4), [2, 2, 2], 7)])
def test_direct26_halo_over_gloo(world, n, grid, peers):
    """csrc/halo26.cu: 2x1x1 (one peer, two directions each way plus self images in y and z), 1x2x2 and 2x2x2 (7 peers,
    26 directions, every peer reached through several of them)"""
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_h26, args=(world, 29563 + world, out, n), nprocs=world, join=True)
    assert len(out) == world
    for r in range(world):
        same_ghosts, same_forward, rev_err, nghost, npeers, inside, total, nmoved, g = out[r]
        assert g == grid and npeers == peers
        assert nghost > 0 and nmoved > 0
        assert same_ghosts, r           # ghost multiset bit-identical to CommBrick's staged borders
        assert same_forward, r
        assert rev_err < 1e-12, (r, rev_err)
        assert inside and total == 4 * n[0] * n[1] * n[2]      # every atom owned exactly once, inside its owner's sub-box
