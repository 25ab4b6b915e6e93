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


if __name__ == "__main__":
    make("lj_n5", O.IN_LJ.format(nx=5, ny=5, nz=5), "lj/cut", 5, 2.5)
    pot = os.path.join(O.POT_DIR, "Cu_u3.eam")
    make("eam_n4", O.IN_EAM.format(nx=4, ny=4, nz=4, style="eam", coeff="1 1 " + pot), "eam", 4, 4.9)
    pot = os.path.join(O.POT_DIR, "Cu_mishin1.eam.alloy")
    make("eamalloy_n4", O.IN_EAM.format(nx=4, ny=4, nz=4, style="eam/alloy", coeff="* * %s Cu" % pot), "eam/alloy", 4, 5.5)
