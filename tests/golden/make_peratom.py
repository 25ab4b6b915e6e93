"""Generates tests/golden/peratom.npz from the UNMODIFIED reference (oracle/_ref/liblammps_ref.so):
`compute pe/atom` and `compute stress/atom NULL [virial]` (src/compute_pe_atom.cpp,
src/compute_stress_atom.cpp — the consumers of Pair::ev_tally's eflag_atom / vflag_atom branches,
src/pair.cpp:1036-1040,1078-1097, and of Bond::ev_tally's, src/bond.cpp:186-222) at the SAME
states the other fixtures hold (lj_n5 / eam_n4 / eamalloy_n4: 25 steps then `run 0`;
chain: data.chain at step 0, every 8th tag), in tag order.  The generator asserts that the
positions equal the ones already committed, so the fixtures stay one consistent set.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import oracle as O  # noqa: E402
from make_golden import IN_CHAIN  # noqa: E402

PER = """
compute pea all pe/atom
compute sva all stress/atom NULL virial
compute sa all stress/atom NULL
compute rpe all reduce sum c_pea
compute rsv all reduce sum c_sva[1] c_sva[2] c_sva[3] c_sva[4] c_sva[5] c_sva[6]
compute rs all reduce sum c_sa[1] c_sa[2] c_sa[3]
thermo_style custom step pe c_rpe c_rsv[1] c_rsv[4] c_rs[1] press vol
thermo_modify norm no
"""


def gather_compute(R, cid, count):
    out = np.zeros(R.natoms() * count)
    R.L.lammps_gather(R.h, cid.encode(), 1, count, out.ctypes.data_as(O.C.c_void_p))
    return out.reshape(-1, count) if count > 1 else out


def grab(R, out, name, sel=slice(None)):
    out[name + "_pe"] = gather_compute(R, "c_pea", 1)[sel]
    out[name + "_stress_virial"] = gather_compute(R, "c_sva", 6)[sel]
    out[name + "_stress"] = gather_compute(R, "c_sa", 6)[sel]
    # whole-system sums (all atoms, also when sel thins the per-atom rows)
    out[name + "_pe_sum"] = gather_compute(R, "c_pea", 1).sum()
    out[name + "_stress_virial_sum"] = gather_compute(R, "c_sva", 6).sum(axis=0)
    out[name + "_thermo_pe"] = R.thermo("pe")
    out[name + "_thermo_press"] = R.thermo("press")
    out[name + "_vol"] = R.thermo("vol")
    print(name, out[name + "_pe"].shape, out[name + "_pe_sum"], out[name + "_thermo_pe"])


def atomic(out, name, script):
    g = np.load(os.path.join(HERE, name + ".npz"))
    R = O.Reference()
    R.command(script)
    R.command(PER)
    R.command("run 25")
    R.command("run 0")
    assert np.array_equal(R.gather("x", 3), g["x"]), name + ": state differs from the committed fixture"
    grab(R, out, name)
    R.close()


def main():
    out = {}
    atomic(out, "lj_n5", O.IN_LJ.format(nx=5, ny=5, nz=5))
    pot = os.path.join(O.POT_DIR, "Cu_u3.eam")
    atomic(out, "eam_n4", O.IN_EAM.format(nx=4, ny=4, nz=4, style="eam", coeff="1 1 " + pot))
    pot = os.path.join(O.POT_DIR, "Cu_mishin1.eam.alloy")
    atomic(out, "eamalloy_n4", O.IN_EAM.format(nx=4, ny=4, nz=4, style="eam/alloy", coeff="* * %s Cu" % pot))
    # bench/in.chain without fix langevin (see make_golden.make_chain), step 0
    R = O.Reference(args=("-log", "none", "-echo", "none", "-nocite"))
    R.command(IN_CHAIN.format(data=os.path.join(os.path.dirname(O.POT_DIR), "data", "data.chain")))
    R.command(PER)
    R.command("run 0")
    grab(R, out, "chain", slice(None, None, 8))
    R.close()
    np.savez_compressed(os.path.join(HERE, "peratom.npz"), **out)


if __name__ == "__main__":
    main()
