"""CPU: the C oracle against the UNMODIFIED reference run live (oracle/_ref/liblammps_ref.so; skipped
where it was not built).  Starts from the lattice, so it also covers create_atoms + velocity."""
import os

import numpy as np
import pytest

import common as Cm
from conftest import have_ref

pytestmark = pytest.mark.skipif(not have_ref(), reason="oracle/_ref not built (needs /root/reference)")
KEYS = ("temp", "pe", "ke", "press")


def _compare(o, R, style, steps, dt, tol_f=1e-11, tol_t=1e-10):
    o.setup()
    R.command("thermo_modify norm no\nrun 0")
    a = o.atoms()
    nl = R.neighlist_by_tag(style)
    ti = np.array(sorted(nl)); nn = np.array([len(nl[t]) for t in ti]); flat = np.concatenate([nl[t] for t in ti])
    nno, flo = o.neigh_tags()
    assert np.array_equal(Cm.pairs_from_lists(ti, nn, flat), Cm.pairs_from_lists(a["tag"], nno, flo))
    for k in KEYS:
        assert abs(o.thermo()[k] - R.thermo(k)) <= tol_t * abs(R.thermo(k)), k
    R.command("run %d" % steps)
    t = o.run(steps, dt, 0)[-1]
    for k in KEYS:
        assert abs(t[k] - R.thermo(k)) <= tol_t * abs(R.thermo(k)), (k, t[k], R.thermo(k))
    a = o.atoms()
    order = np.argsort(a["tag"])
    fr = R.gather("f", 3)
    assert np.abs(a["f"][order] - fr).max() <= tol_f * np.abs(fr).max()
    assert np.abs(a["x"][order] - R.gather("x", 3)).max() <= 1e-11


def test_lj(pkg, oracle_mod):
    O = oracle_mod
    s = Cm.lj_system(pkg, 5)
    T = pkg.b200md.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    o = Cm.make_oracle(O, pkg, s, "lj", T, 20, 0, 0)
    R = O.Reference()
    R.command(O.IN_LJ.format(nx=5, ny=5, nz=5))
    _compare(o, R, "lj/cut", 100, 0.005)
    R.close(); o.close()


@pytest.mark.parametrize("style,pot,elems", [("eam", "Cu_u3.eam", None), ("eam/alloy", "Cu_mishin1.eam.alloy", ["Cu"])])
def test_eam(pkg, oracle_mod, style, pot, elems):
    O = oracle_mod
    path = os.path.join(O.POT_DIR, pot)
    T = pkg.eam_host.tables_from_funcfl([path]) if elems is None else pkg.eam_host.tables_from_setfl(path, elems)
    s = Cm.eam_system(pkg, 4, T["mass"][1])
    o = Cm.make_oracle(O, pkg, s, "eam", T, 1, 5, 1)
    R = O.Reference()
    coeff = "1 1 " + path if elems is None else "* * %s Cu" % path
    R.command(O.IN_EAM.format(nx=4, ny=4, nz=4, style=style, coeff=coeff))
    _compare(o, R, style, 100, 0.005)
    R.close(); o.close()
