"""Shared set-up for the parity tests: the same seeded input built three ways — the CUDA
path (b200md Context), the C oracle, and the unmodified reference (when oracle/_ref exists)."""
import math
import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def lj_system(pkg, n):
    """bench/in.lj geometry with n^3 fcc cells: box, atoms with final velocities (host numpy)."""
    B = pkg.b200md
    scale = math.pow(4.0 / 0.8442, 1.0 / 3.0)
    boxlo, boxhi = [0.0] * 3, [n * scale] * 3
    x, v, tag = B.lattice_fcc(scale, n, n, n, boxlo, [1e30] * 3, 87287, 1.0)
    return dict(boxlo=boxlo, boxhi=boxhi, x=x, v=v, tag=tag, type=np.ones(len(tag), np.int32),
                mass=np.array([0.0, 1.0]), units=B.UNITS["lj"], temp=1.44, dt=0.005, skin=0.3)


def eam_system(pkg, n, mass):
    B = pkg.b200md
    scale = 3.615
    boxlo, boxhi = [0.0] * 3, [n * scale] * 3
    x, v, tag = B.lattice_fcc(scale, n, n, n, boxlo, [1e30] * 3, 376847, mass)
    return dict(boxlo=boxlo, boxhi=boxhi, x=x, v=v, tag=tag, type=np.ones(len(tag), np.int32),
                mass=np.array([0.0, mass]), units=B.UNITS["metal"], temp=1600.0, dt=0.005, skin=1.0)


def make_oracle(O, pkg, sysd, pair, tables, every, delay, check, half=True):
    B = pkg.b200md
    o = O.Oracle(sysd["boxlo"], sysd["boxhi"], sysd["x"], sysd["v"], sysd["type"], sysd["tag"], sysd["mass"], sysd["units"])
    o.velocity_finalize(sysd["temp"])
    if pair == "lj":
        o.pair_lj(tables)
        cutsq = tables["cutsq"]
    else:
        o.pair_eam(tables)
        cutsq = np.zeros((2, 2)); cutsq[1, 1] = tables["cutforcesq"]
    cn, cmax = B.cutneighsq_table(1, cutsq, sysd["skin"])
    o.neigh_config(cn, cmax, sysd["skin"], every, delay, check, half)
    return o


def pairs_from_lists(tags_i, numneigh, flat):
    """unordered (min,max) tag pairs of a half list as a sorted int64 key array"""
    i = np.repeat(tags_i.astype(np.int64), numneigh)
    j = flat.astype(np.int64)
    lo, hi = np.minimum(i, j), np.maximum(i, j)
    return np.sort(lo * (1 << 32) + hi)


def full_sets(tags_i, numneigh, flat):
    """per-atom sorted neighbor tag multisets of a full list, ordered by tag_i: (tag_i, tag_j) keys"""
    i = np.repeat(tags_i.astype(np.int64), numneigh)
    return np.sort(i * (1 << 32) + flat.astype(np.int64))
