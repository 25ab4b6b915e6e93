"""CPU: host-side logic of libb200md.so (csrc/host.cpp) — decomposition, swap schedule, bin geometry,
lattice + velocity generator — and the Python host mirror (LJ tables, EAM tables)."""
import math
import os

import numpy as np
import pytest

import common as Cm
from conftest import have_ref


def test_procgrid_matches_reference_logs(pkg):
    B = pkg.b200md
    # bench/log.6Oct16.lj.fixed.icc.4: "1 by 2 by 2 MPI processor grid" (cubic box);
    # bench/log.6Oct16.lj.scaled.icc.4: "2 by 2 by 1" for the 2x2x1-scaled box
    assert B.procgrid(4, [33.6, 33.6, 33.6]) == [1, 2, 2]
    assert B.procgrid(4, [67.2, 67.2, 33.6]) == [2, 2, 1]
    assert B.procgrid(8, [1, 1, 1]) == [2, 2, 2]
    assert B.procgrid(2, [2, 1, 1]) == [2, 1, 1]
    assert B.procgrid(1, [1, 1, 1]) == [1, 1, 1]
    g = [2, 2, 2]
    locs = [tuple(B.rank2loc(r, g)) for r in range(8)]
    assert len(set(locs)) == 8 and locs[1] == (0, 0, 1) and locs[4] == (1, 0, 0)


def test_local_box_tiles_domain(pkg):
    B = pkg.b200md
    lo, hi, g = [0.0, -1.0, 2.0], [10.0, 5.0, 9.0], [2, 3, 1]
    vol = 0.0
    for r in range(6):
        sl, sh = B.local_box(lo, hi, g, B.rank2loc(r, g))
        vol += np.prod(np.array(sh) - np.array(sl))
        if B.rank2loc(r, g)[0] == 1:
            assert sh[0] == hi[0]  # last proc ends exactly at boxhi (domain.cpp:347-348)
    assert abs(vol - 10 * 6 * 7) < 1e-9


def test_swap_schedule_single_rank(pkg):
    B = pkg.b200md
    s = B.comm_schedule([0, 0, 0], [33.6, 33.6, 33.6], [1, 1, 1], [1, 1, 1], [0, 0, 0], 2.8)
    assert s.nswap == 6 and list(s.maxneed) == [1, 1, 1]
    for k in range(6):
        assert s.sendproc[k] == 0 and s.recvproc[k] == 0 and s.sendflag[k] == 1
        d = s.dim[k]
        assert list(s.pbc[k])[d] == (1 if k % 2 == 0 else -1)  # both edges are periodic edges
    assert s.slabhi[0] == pytest.approx(2.8) and s.slablo[0] < -1e19
    assert s.slablo[1] == pytest.approx(33.6 - 2.8) and s.slabhi[1] > 1e19


def test_swap_schedule_two_ranks(pkg):
    B = pkg.b200md
    for me in (0, 1):
        s = B.comm_schedule([0, 0, 0], [20, 10, 10], [1, 1, 1], [2, 1, 1], B.rank2loc(me, [2, 1, 1]), 2.5)
        assert s.nswap == 6
        assert s.sendproc[0] == 1 - me and s.recvproc[0] == 1 - me and s.sendproc[2] == me
        # shift only where the swap crosses the periodic edge (comm_brick.cpp:436-468)
        assert list(s.pbc[0])[0] == (1 if me == 0 else 0)
        assert list(s.pbc[1])[0] == (-1 if me == 1 else 0)
    s = B.comm_schedule([0, 0, 0], [4, 10, 10], [1, 1, 1], [1, 1, 1], [0, 0, 0], 5.0)
    assert list(s.maxneed) == [2, 1, 1] and s.nswap == 8


def test_bin_geometry_matches_oracle(pkg, oracle_mod):
    B = pkg.b200md
    L = 20 * (4.0 / 0.8442) ** (1.0 / 3.0)
    g = B.bin_setup([0, 0, 0], [L] * 3, [0, 0, 0], [L] * 3, 2.8, 2.8)
    # int(33.59/1.4) = 23 bins of 1.4605 (the reference's log line "bins = 24 24 24" prints ceil())
    assert (g.nbinx, g.nbiny, g.nbinz) == (23, 23, 23) and (g.sx, g.sy, g.sz) == (2, 2, 2)
    s = Cm.lj_system(pkg, 20)
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    o = Cm.make_oracle(oracle_mod, pkg, s, "lj", T, 20, 0, 0, half=False)
    ints, dbl = o.bin_geometry()
    assert ints[:9] == [g.nbinx, g.nbiny, g.nbinz, g.mbinx, g.mbiny, g.mbinz, g.mbinxlo, g.mbinylo, g.mbinzlo]
    assert dbl == [g.binsizex, g.binsizey, g.binsizez, g.bininvx, g.bininvy, g.bininvz]
    assert ints[9] == 125  # full stencil
    o.close()


def test_lj_tables(pkg):
    B = pkg.b200md
    T = B.lj_cut_tables(2, {(1, 1): (1.0, 1.0), (2, 2): (0.5, 1.2, 3.0)}, 2.5, mix="arithmetic", shift=True)
    assert T["lj1"][1, 1] == 48.0 and T["lj2"][1, 1] == 24.0 and T["cutsq"][1, 1] == 6.25
    e, s, c = math.sqrt(0.5), 1.1, 2.75   # cutoff mixed like sigma: mix_distance(2.5, 3.0), src/pair_lj_cut.cpp:506
    assert T["lj3"][1, 2] == pytest.approx(4 * e * s ** 12) and T["lj3"][2, 1] == T["lj3"][1, 2]
    assert T["offset"][1, 2] == pytest.approx(4 * e * ((s / c) ** 12 - (s / c) ** 6))
    with pytest.raises(B.B200MDError, match="All pair coeffs are not set"):
        B.lj_cut_tables(2, {(1, 1): (1.0, 1.0)}, 2.5)
    cn, cmax = B.cutneighsq_table(1, np.array([[0, 0], [0, 6.25]]), 0.3)
    assert cn[1, 1] == (2.5 + 0.3) ** 2 and cmax == 2.8


def test_lattice_counts_and_partition(pkg):
    B = pkg.b200md
    scale = 3.615
    x, v, tag = B.lattice_fcc(scale, 5, 4, 3, [0, 0, 0], [1e9] * 3, 376847, 63.55)
    assert len(tag) == 4 * 60 and sorted(tag) == list(range(1, 241))
    tot = 0
    for r in range(4):
        lo, hi = B.local_box([0, 0, 0], [5 * scale, 4 * scale, 3 * scale], [2, 2, 1], B.rank2loc(r, [2, 2, 1]))
        xs, vs, ts = B.lattice_fcc(scale, 5, 4, 3, lo, hi, 376847, 63.55)
        tot += len(ts)
        # loop geom: a rank's velocities are the same draws the single-rank run makes
        idx = np.searchsorted(np.sort(tag), ts)
        assert np.array_equal(v[np.argsort(tag)][idx], vs)
    assert tot == 240


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built")
def test_lattice_and_velocity_bit_exact_vs_reference(pkg, oracle_mod):
    """create_atoms + velocity loop geom (+ zero momentum + rescale through the oracle) vs the reference."""
    O = oracle_mod
    s = Cm.lj_system(pkg, 6)
    T = pkg.b200md.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    o = Cm.make_oracle(O, pkg, s, "lj", T, 20, 0, 0)
    R = O.Reference()
    R.command(O.IN_LJ.format(nx=6, ny=6, nz=6))
    xr, vr = R.gather("x", 3), R.gather("v", 3)
    a = o.atoms()
    assert np.array_equal(a["x"], xr)  # tags = creation order
    assert np.array_equal(a["v"], vr)
    R.close(); o.close()


def test_eam_tables_shapes(pkg, oracle_mod):
    pot = os.path.join(oracle_mod.POT_DIR, "Cu_u3.eam")
    if not os.path.exists(pot):
        pytest.skip("potential file not present")
    tb = pkg.eam_host.tables_from_funcfl([pot])
    assert tb["nr"] == 499 and tb["nrho"] == 499 and tb["cutforcesq"] == pytest.approx(4.95 ** 2)
    assert tb["frho_spline"].shape == (2, 500, 7) and tb["z2r_spline"].shape == (1, 500, 7)
    sp = tb["rhor_spline"][0]
    # interpolate(): derivative rows are the value cubic's coefficients over delta (pair_eam.cpp:792-796)
    assert np.allclose(sp[1:, 2], sp[1:, 5] * tb["rdr"], rtol=1e-15)
    assert np.allclose(sp[1:, 0], 3 * sp[1:, 3] * tb["rdr"], rtol=1e-15)


@pytest.mark.parametrize("periodic", [(1, 1, 1), (1, 0, 1)])
def test_one_rank_product_set_is_the_staged_ghost_set(pkg, periodic):
    """csrc/comm.cu comm_borders_self26: on one rank the ghosts of CommBrick::borders (comm_brick.cpp:771-979, three
    stages, each forwarding the ghosts of the stages before it) are the product set over the dimensions of the
    per-dimension send slabs — an owned atom has an image in direction (dx,dy,dz) != 0 iff it lies in the slab of
    every non-zero component — with the same positions bit for bit (one addition of pbc*prd per crossed dimension,
    AtomVec::pack_border).  Both restated in numpy on the library's own swap schedule."""
    B = pkg.b200md
    rng = np.random.default_rng(7)
    boxlo, boxhi = [0.0, -3.0, 1.0], [21.0, 17.5, 24.0]
    prd = np.array(boxhi) - np.array(boxlo)
    cut = 5.95
    n = 4000
    x = np.array(boxlo) + rng.random((n, 3)) * prd
    tag = np.arange(1, n + 1)
    s = B.comm_schedule(boxlo, boxhi, list(periodic), [1, 1, 1], [0, 0, 0], cut)
    assert max(s.maxneed) <= 1
    # staged
    X, T = x.copy(), tag.copy()
    iswap = 0
    for dim in range(3):
        nlast = 0
        for ineed in range(2 * s.maxneed[dim]):
            if ineed % 2 == 0:
                nfirst, nlast = nlast, len(T)
            sel = (np.nonzero((X[nfirst:nlast, dim] >= s.slablo[iswap]) & (X[nfirst:nlast, dim] <= s.slabhi[iswap]))[0] + nfirst
                   if s.sendflag[iswap] else np.empty(0, np.int64))
            buf = X[sel].copy()
            for d in range(3):
                if s.pbc[iswap][d]:
                    buf[:, d] = buf[:, d] + s.pbc[iswap][d] * prd[d]
            X = np.concatenate([X, buf]); T = np.concatenate([T, T[sel]])
            iswap += 1
    staged = sorted(zip(T[n:].tolist(), map(tuple, X[n:].tolist())))
    # product set
    L, R, pL, pR = [None] * 3, [None] * 3, [0] * 3, [0] * 3
    iswap = 0
    for dim in range(3):
        L[dim] = np.zeros(n, bool); R[dim] = np.zeros(n, bool)
        if s.maxneed[dim] == 1:
            if s.sendflag[iswap]:
                L[dim] = (x[:, dim] >= s.slablo[iswap]) & (x[:, dim] <= s.slabhi[iswap]); pL[dim] = s.pbc[iswap][dim]
            if s.sendflag[iswap + 1]:
                R[dim] = (x[:, dim] >= s.slablo[iswap + 1]) & (x[:, dim] <= s.slabhi[iswap + 1]); pR[dim] = s.pbc[iswap + 1][dim]
            iswap += 2
    prod = []
    for dz in (-1, 0, 1):
        for dy in (-1, 0, 1):
            for dx in (-1, 0, 1):
                comp = (dx, dy, dz)
                if comp == (0, 0, 0):
                    continue
                ok = np.ones(n, bool)
                for k in range(3):
                    if comp[k] < 0: ok &= L[k]
                    elif comp[k] > 0: ok &= R[k]
                img = x[ok].copy()
                for k in range(3):
                    p = pL[k] if comp[k] < 0 else (pR[k] if comp[k] > 0 else 0)
                    if p:
                        img[:, k] = img[:, k] + p * prd[k]
                prod += list(zip(tag[ok].tolist(), map(tuple, img.tolist())))
    assert len(staged) > 0 and sorted(prod) == staged
