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
