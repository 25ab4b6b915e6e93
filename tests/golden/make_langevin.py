"""Generates tests/golden/langevin.npz from the UNMODIFIED reference (oracle/_ref/liblammps_ref.so):
`fix langevin` (src/fix_langevin.cpp) on top of the two states the other fixtures use.

chain_*  bench/in.chain exactly as the reference ships it (fix nve + fix langevin 1.0 1.0 10.0
         904297): forces after the set-up (`run 0`; they already hold the thermostat's drag and
         random force, FixLangevin::setup), thermo every 10 steps of the following 100
         (`run 10 pre no post no`, i.e. one uninterrupted run), x / v of every 8th tag at the end.
rep212_* in.chain.scaled (replicate 2 1 2, 128 000 atoms) with the thermostat: thermo after `run 20`
         (the draws follow the creation order `replicate` leaves, then Atom::sort).
ljsort_* the lj_n5 state, fix langevin 1.0 1.0 1.0 7771, ONE `run 1100` from timestep 0: the reference
         re-sorts its atoms at the first reneighboring at or after step 1000 (Atom::sort, verlet.cpp:285),
         which changes the order the RanMars draws are handed out in; x, v, thermo at the end.
lj_*     the lj_n5 state with a temperature ramp and a scale factor
         (fix langevin 1.0 2.5 0.5 48271 scale 1 1.5), one `run 40`; x, v, f of all atoms.  The
         RanMars draws follow the reference's local atom order, so that order (tags) is stored.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, HERE)
from oracle import oracle as O  # noqa: E402
from make_golden import IN_CHAIN  # noqa: E402

KW = ("temp", "pe", "ke", "press", "evdwl", "emol")
LJ_FIX = "fix 2 all langevin 1.0 2.5 0.5 48271 scale 1 1.5"
CHAIN_FIX = "fix 2 all langevin 1.0 1.0 10.0 904297"


def main():
    out = {}
    R = O.Reference(args=("-log", "none", "-echo", "none", "-nocite"))
    R.command(IN_CHAIN.format(data=os.path.join(os.path.dirname(O.POT_DIR), "data", "data.chain")))
    R.command(CHAIN_FIX)
    R.command("thermo_style custom step temp pe ke press evdwl emol")
    R.command("thermo_modify norm no")
    R.command("run 0")
    f = R.gather("f", 3)
    out["chain_f_every8"] = f[::8].copy()
    out["chain_f_abs_sum"] = np.abs(f).sum(axis=0)
    out["chain_thermo0"] = np.array([R.thermo(k) for k in KW])
    series = []
    for _ in range(10):
        R.command("run 10 pre no post no")
        series.append([R.thermo(k) for k in KW])
    out["chain_series"] = np.array(series)
    out["chain_x_every8"] = R.gather("x", 3)[::8].copy()
    out["chain_v_every8"] = R.gather("v", 3)[::8].copy()
    R.close()

    # in.chain.scaled with -var x 2 -var y 1 -var z 2 (128 000 atoms) + the thermostat: thermo only
    R = O.Reference(args=("-log", "none", "-echo", "none", "-nocite"))
    R.command(IN_CHAIN.format(data=os.path.join(os.path.dirname(O.POT_DIR), "data", "data.chain")).replace(
        "neighbor 0.4 bin", "replicate 2 1 2\nneighbor 0.4 bin"))
    R.command(CHAIN_FIX)
    R.command("thermo_style custom step temp pe ke press evdwl emol")
    R.command("thermo_modify norm no")
    R.command("run 20")
    out["rep212_thermo20"] = np.array([R.thermo(k) for k in KW])
    R.close()

    R = O.Reference()
    R.command(O.IN_LJ.format(nx=5, ny=5, nz=5))
    R.command("thermo_style custom step temp pe ke press")
    R.command("thermo_modify norm no")
    R.command("run 25")
    R.command("run 0")
    nl = R.setting("nlocal")
    out["lj_local_tags"] = R.local_ids()[:nl].astype(np.int32)
    out["lj_x0"] = R.gather("x", 3); out["lj_v0"] = R.gather("v", 3)
    R.command(LJ_FIX)
    R.command("run 40")
    out["lj_x"] = R.gather("x", 3); out["lj_v"] = R.gather("v", 3); out["lj_f"] = R.gather("f", 3)
    out["lj_image"] = R.gather("image", 1, np.int32)
    out["lj_thermo"] = np.array([R.thermo(k) for k in KW[:4]])
    R.close()
    # a run that crosses the reference's re-sort step (Atom::sort every 1000 steps, verlet.cpp:285):
    # after it the RanMars draws follow the NEW local order
    R = O.Reference()
    R.command(O.IN_LJ.format(nx=5, ny=5, nz=5))
    R.command("thermo_style custom step temp pe ke press")
    R.command("thermo_modify norm no")
    R.command("run 25")
    R.command("reset_timestep 0")
    R.command("run 0")
    nl = R.setting("nlocal")
    out["ljsort_local_tags"] = R.local_ids()[:nl].astype(np.int32)
    out["ljsort_x0"] = R.gather("x", 3); out["ljsort_v0"] = R.gather("v", 3)
    R.command("fix 2 all langevin 1.0 1.0 1.0 7771")
    R.command("run 1100")
    out["ljsort_x"] = R.gather("x", 3); out["ljsort_v"] = R.gather("v", 3)
    out["ljsort_thermo"] = np.array([R.thermo(k) for k in KW[:4]])
    out["ljsort_local_tags_end"] = R.local_ids()[:R.setting("nlocal")].astype(np.int32)
    R.close()
    np.savez_compressed(os.path.join(HERE, "langevin.npz"), **out)
    print("chain", out["chain_thermo0"], out["chain_series"][-1])
    print("lj", out["lj_thermo"])


if __name__ == "__main__":
    main()
