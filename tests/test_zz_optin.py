"""Opt-in device paths that are off by default (they sort last on purpose: a failure here must not
hide the parity results of the default paths).

B200MD_FUSE_NVE=1: FixNVE::final_integrate of step n and initial_integrate of step n+1
(src/fix_nve.cpp:64-141) in one pass over v / f (csrc/nve.cu: k_nve_final_initial).  Every rounding
stays where the two separate kernels have it, so on the full list (no atomics anywhere) the whole
trajectory must be BIT-IDENTICAL, rebuild decisions and thermo included."""
import os

import numpy as np
import pytest

import common as Cm
from test_oracle_golden import CASES


def trajectory(pkg, oracle_mod, name, kind, fused):
    from test_gpu_parity import make_ctx
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    old = os.environ.get("B200MD_FUSE_NVE")
    os.environ["B200MD_FUSE_NVE"] = "1" if fused else "0"      # read by b200md_create
    try:
        ctx, s = make_ctx(pkg, oracle_mod, g, kind, B.LIST_FULL, 2)
    finally:
        if old is None:
            del os.environ["B200MD_FUSE_NVE"]
        else:
            os.environ["B200MD_FUSE_NVE"] = old
    ctx.set_option("graph", "off")           # the launch counts below compare the host-driven loops
    ctx.setup()
    th = ctx.run(60, s["dt"], 25)            # thermo steps at 25, 50, 60 break the fusion; rebuilds fall in between
    th += ctx.run(7, s["dt"], 0)
    a = ctx.get_atoms(("x", "v", "f", "tag"))
    st = ctx.neigh_stats()
    launches = ctx.timing()["launches"]
    ctx.close()
    o = np.argsort(a["tag"])
    return a["x"][o], a["v"][o], a["f"][o], th, st, launches


@pytest.mark.gpu
@pytest.mark.parametrize("name,kind", CASES)
def test_fused_nve_is_bit_identical(pkg, oracle_mod, name, kind):
    x0, v0, f0, th0, st0, l0 = trajectory(pkg, oracle_mod, name, kind, False)
    x1, v1, f1, th1, st1, l1 = trajectory(pkg, oracle_mod, name, kind, True)
    assert np.array_equal(x0, x1) and np.array_equal(v0, v1) and np.array_equal(f0, f1)
    assert st0["nbuilds"] == st1["nbuilds"] and st0["ndanger"] == st1["ndanger"]
    for a, b in zip(th0, th1):
        assert a["step"] == b["step"]
        for k in ("temp", "pe", "ke", "press"):
            assert a[k] == b[k], (k, a[k], b[k])
    assert l1 < l0                            # one launch fewer per fused step pair


@pytest.mark.gpu
@pytest.mark.parametrize("half", [True, False])
def test_cuda_langevin_across_resort(pkg, half):
    """fix langevin over ONE run of 1100 steps: at the first rebuild at or after step 1000 the reference
    re-sorts its atoms (Atom::sort, src/verlet.cpp:285) and hands out the RanMars draws in the new order;
    the device re-ranks its stream index at the same rebuild (csrc/langevin.cu: langevin_resort).
    Golden: tests/golden/langevin.npz ljsort_* from the unmodified reference.  Rounding differences
    grow over 1100 steps (5.5 tau) of a melt, hence 2e-4; a stream left in the old order is off by 3e-3 .. 8e-2
    (measured with the oracle)."""
    import test_langevin as TL
    G = TL.golden()
    B = pkg.b200md
    s = TL.ljsort_inputs(pkg, G)
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.3)
    perm = pkg.inputs.atom_sort_order(s["x"], s["boxlo"], s["boxhi"], cmax)
    ctx = B.Context(0)
    ctx.set_units("lj"); ctx.set_box(s["boxlo"], s["boxhi"])
    ctx.set_atoms(s["x"][perm], s["v"][perm], s["type"][perm], s["tag"][perm], s["mass"])
    ctx.pair_lj_cut_config(*[T[k] for k in ("cutsq", "lj1", "lj2", "lj3", "lj4", "offset")])
    ctx.neigh_config(cn, cmax, 0.3, every=20, delay=0, check=0, list_kind=B.LIST_HALF_NEWTON if half else B.LIST_FULL)
    ctx.fix_langevin(1.0, 1.0, 1.0, 7771, 0.005)
    ctx.setup()
    t = ctx.run(1100, 0.005, 0)[-1]
    a = ctx.get_atoms(("x", "v", "tag"))
    TL.check_ljsort(G, s, a["tag"], a["x"], a["v"], t, 2e-4, 2e-4)
    ctx.close()
