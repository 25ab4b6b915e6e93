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
