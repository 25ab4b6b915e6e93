"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/liblammps_ref.so,
built by oracle/build_ref.py from /root/reference).  Run here, where the reference exists;
the fixtures travel to the GPU box.

Each fixture: a disordered state (25 steps of the bench input at a small size, then `run 0`
so positions are wrapped and lists fresh), with tag-ordered x, v, f at full precision, the
half/newton pair list and the full list as tag keys, thermo at that state, and the thermo
series of the following 100 steps (every 10).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle as O  # noqa: E402

KW = ("temp", "pe", "ke", "press")


def keys(nl):
    ti = np.array(sorted(nl), dtype=np.int64)
    i = np.repeat(ti, [len(nl[t]) for t in ti])
    j = np.concatenate([nl[t] for t in ti]).astype(np.int64)
    return i, j


def make(name, script, style, n, cutoff):
    R = O.Reference()
    R.command(script)
    R.command("compute c all coord/atom cutoff %g" % cutoff)
    R.command("compute red all reduce sum c_c")
    R.command("thermo_style custom step temp pe ke press c_red")
    R.command("thermo_modify norm no")
    R.command("run 25")
    R.command("run 0")
    out = dict(n=n, natoms=R.natoms())
    out["x"] = R.gather("x", 3); out["v"] = R.gather("v", 3); out["f"] = R.gather("f", 3)
    out["image"] = R.gather("image", 1, np.int32)
    out["thermo0"] = np.array([R.thermo(k) for k in KW])
    i, j = keys(R.neighlist_by_tag(style))
    lo, hi = np.minimum(i, j), np.maximum(i, j)
    o = np.lexsort((hi, lo))
    out["half_lo"] = lo[o].astype(np.int32); out["half_hi"] = hi[o].astype(np.int32)
    i, j = keys(R.full_neighlist_by_tag("c"))
    o = np.lexsort((j, i))
    out["full_i"] = i[o].astype(np.int32); out["full_j"] = j[o].astype(np.int32)
    series = []
    for _ in range(10):
        R.command("run 10")
        series.append([R.thermo(k) for k in KW])
    out["thermo_series"] = np.array(series)
    R.close()
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, out["natoms"], len(out["half_lo"]), len(out["full_i"]), out["thermo0"], out["thermo_series"][-1])


IN_CHAIN = """
units lj
atom_style bond
atom_modify map array
special_bonds fene
read_data {data}
neighbor 0.4 bin
neigh_modify every 1 delay 1
bond_style fene
bond_coeff 1 30.0 1.5 1.0 1.0
pair_style lj/cut 1.12
pair_modify shift yes
pair_coeff 1 1 1.0 1.0 1.12
fix 1 all nve
timestep 0.012
"""


def pair_hash(lo, hi):
    """two independent 64-bit digests of the set of (lo,hi) tag pairs (order-free)"""
    k = lo.astype(np.uint64) * np.uint64(1 << 32) + hi.astype(np.uint64)
    m = k * np.uint64(0x9E3779B97F4A7C15)
    m ^= m >> np.uint64(29)
    m *= np.uint64(0xBF58476D1CE4E5B9)
    return np.array([np.bitwise_xor.reduce(m), m.sum(dtype=np.uint64)], dtype=np.uint64)


def make_chain():
    """bench/in.chain WITHOUT `fix langevin` (its RanMars stream is drawn in local-index order,
    src/fix_langevin.cpp:85, which no other ordering of the atoms can replay): state of
    data.chain itself.  32 000 atoms are too many for a fixture, so: forces of every 8th tag,
    |f| checksums, numneigh per tag, digests of the half pair list, thermo (incl. evdwl / emol)
    at step 0 and after 10 x `run 10`."""
    # (no "-screen none": this fork's AtomVec border debug print dereferences the screen handle)
    R = O.Reference(args=("-log", "none", "-echo", "none", "-nocite"))
    R.command(IN_CHAIN.format(data=os.path.join(os.path.dirname(O.POT_DIR), "data", "data.chain")))
    kw = ("temp", "pe", "ke", "press", "evdwl", "emol")
    R.command("thermo_style custom step temp pe ke press evdwl emol")
    R.command("thermo_modify norm no")
    R.command("run 0")
    out = dict(natoms=R.natoms())
    f = R.gather("f", 3)
    out["f_every8"] = f[::8].copy()
    out["f_abs_sum"] = np.abs(f).sum(axis=0)
    out["f_sq_sum"] = (f * f).sum()
    out["thermo0"] = np.array([R.thermo(k) for k in kw])
    nl = R.neighlist_by_tag("lj/cut")
    i, j = keys(nl)
    lo, hi = np.minimum(i, j), np.maximum(i, j)
    out["npairs"] = len(lo)
    out["pair_digest"] = pair_hash(lo, hi)
    cnt = np.zeros(out["natoms"] + 1, np.int32)
    np.add.at(cnt, lo, 1); np.add.at(cnt, hi, 1)
    out["pairs_per_tag"] = cnt.astype(np.uint8)          # appearances of each tag in the pair set
    series = []
    for _ in range(10):
        R.command("run 10")
        series.append([R.thermo(k) for k in kw])
    out["thermo_series"] = np.array(series)
    R.close()
    # in.chain.scaled with -var x 2 -var y 1 -var z 2 (128 000 atoms): thermo only
    R = O.Reference(args=("-log", "none", "-echo", "none", "-nocite"))
    R.command(IN_CHAIN.format(data=os.path.join(os.path.dirname(O.POT_DIR), "data", "data.chain")).replace(
        "neighbor 0.4 bin", "replicate 2 1 2\nneighbor 0.4 bin"))
    R.command("thermo_style custom step temp pe ke press evdwl emol")
    R.command("thermo_modify norm no")
    R.command("run 0")
    out["rep212_thermo0"] = np.array([R.thermo(k) for k in kw])
    series = []
    for _ in range(3):
        R.command("run 10")
        series.append([R.thermo(k) for k in kw])
    out["rep212_series"] = np.array(series)
    R.close()
    np.savez_compressed(os.path.join(HERE, "chain.npz"), **out)
    print("chain", out["natoms"], out["npairs"], out["pair_digest"], out["thermo0"], out["thermo_series"][-1])


IN_LJMIX = """
units lj
atom_style atomic
atom_modify map array
lattice fcc 0.8442
region box block 0 4 0 4 0 4
create_box 2 box
create_atoms 1 box
region left block 0 2 INF INF INF INF
set region left type 2
mass 1 1.0
mass 2 1.5
velocity all create 1.44 87287 loop geom
pair_style lj/cut 2.5
pair_modify mix {mix}
pair_coeff 1 1 1.0 1.0 2.5
pair_coeff 2 2 0.5 0.9 2.0
neighbor 0.3 bin
neigh_modify delay 0 every 20 check no
fix 1 all nve
"""


def make_ljmix():
    """two atom types with DIFFERENT per-type cutoffs and no explicit 1-2 coefficients: epsilon, sigma and
    the cutoff of the mixed pair come from Pair::mix_energy / mix_distance (PairLJCut::init_one,
    src/pair_lj_cut.cpp:501-556), for `pair_modify mix geometric` and `arithmetic`."""
    out = {}
    for mix in ("geometric", "arithmetic"):
        R = O.Reference()
        R.command(IN_LJMIX.format(mix=mix))
        R.command("thermo_style custom step temp pe ke press")
        R.command("thermo_modify norm no")
        R.command("run 25")
        R.command("run 0")
        out[mix + "_x"] = R.gather("x", 3); out[mix + "_v"] = R.gather("v", 3); out[mix + "_f"] = R.gather("f", 3)
        out[mix + "_type"] = R.gather("type", 1, np.int32)
        out[mix + "_thermo0"] = np.array([R.thermo(k) for k in KW])
        i, j = keys(R.neighlist_by_tag("lj/cut"))
        lo, hi = np.minimum(i, j), np.maximum(i, j)
        o = np.lexsort((hi, lo))
        out[mix + "_half_lo"] = lo[o].astype(np.int32); out[mix + "_half_hi"] = hi[o].astype(np.int32)
        R.command("run 20")
        out[mix + "_thermo20"] = np.array([R.thermo(k) for k in KW])
        R.close()
        print("ljmix", mix, len(lo), out[mix + "_thermo0"], out[mix + "_thermo20"])
    np.savez_compressed(os.path.join(HERE, "ljmix.npz"), **out)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "ljmix":
        make_ljmix()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "chain":
        make_chain()
        sys.exit(0)
    make("lj_n5", O.IN_LJ.format(nx=5, ny=5, nz=5), "lj/cut", 5, 2.5)
    pot = os.path.join(O.POT_DIR, "Cu_u3.eam")
    make("eam_n4", O.IN_EAM.format(nx=4, ny=4, nz=4, style="eam", coeff="1 1 " + pot), "eam", 4, 4.9)
    pot = os.path.join(O.POT_DIR, "Cu_mishin1.eam.alloy")
    make("eamalloy_n4", O.IN_EAM.format(nx=4, ny=4, nz=4, style="eam/alloy", coeff="* * %s Cu" % pot), "eam/alloy", 4, 5.5)
