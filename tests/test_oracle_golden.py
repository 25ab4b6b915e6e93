"""CPU: the C oracle (oracle/md_oracle.c) against the golden vectors generated from the
unmodified reference (tests/golden/make_golden.py).  This is what pins the oracle."""
import os

import numpy as np
import pytest

import common as Cm

CASES = [("lj_n5", "lj"), ("eam_n4", "eam"), ("eamalloy_n4", "eam/alloy")]


def tables_for(pkg, O, kind):
    B = pkg.b200md
    if kind == "lj":
        return B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    if kind == "eam":
        return pkg.eam_host.tables_from_funcfl([os.path.join(O.POT_DIR, "Cu_u3.eam")])
    return pkg.eam_host.tables_from_setfl(os.path.join(O.POT_DIR, "Cu_mishin1.eam.alloy"), ["Cu"])


def system_from_golden(pkg, g, kind):
    B = pkg.b200md
    n = int(g["n"])
    if kind == "lj":
        s = Cm.lj_system(pkg, 1)
        scale = (4.0 / 0.8442) ** (1.0 / 3.0)
    else:
        s = Cm.eam_system(pkg, 1, 63.55)
        scale = 3.615
    s["boxlo"], s["boxhi"] = [0.0] * 3, [n * scale] * 3
    nat = int(g["natoms"])
    s.update(x=g["x"], v=g["v"], tag=np.arange(1, nat + 1, dtype=np.int32), type=np.ones(nat, np.int32))
    return s


def build_oracle(O, pkg, g, kind, half):
    B = pkg.b200md
    T = tables_for(pkg, O, kind)
    s = system_from_golden(pkg, g, kind)
    if kind != "lj":
        s["mass"] = np.array([0.0, T["mass"][1]])
    o = O.Oracle(s["boxlo"], s["boxhi"], s["x"], s["v"], s["type"], s["tag"], s["mass"], s["units"])
    if kind == "lj":
        o.pair_lj(T); cutsq = T["cutsq"]; sched = (20, 0, 0)
    else:
        o.pair_eam(T); cutsq = np.zeros((2, 2)); cutsq[1, 1] = T["cutforcesq"]; sched = (1, 5, 1)
    cn, cmax = B.cutneighsq_table(1, cutsq, s["skin"])
    o.neigh_config(cn, cmax, s["skin"], sched[0], sched[1], sched[2], half)
    return o, s


@pytest.mark.parametrize("name,kind", CASES)
def test_half_list_forces_thermo(pkg, oracle_mod, name, kind):
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    o, s = build_oracle(oracle_mod, pkg, g, kind, True)
    o.setup()
    a = o.atoms()
    nn, flat = o.neigh_tags()
    got = Cm.pairs_from_lists(a["tag"], nn, flat)
    want = np.sort(g["half_lo"].astype(np.int64) * (1 << 32) + g["half_hi"].astype(np.int64))
    assert np.array_equal(got, want)  # bit-exact unordered pair set
    order = np.argsort(a["tag"])
    f = a["f"][order]
    fmax = np.abs(g["f"]).max()
    assert np.abs(f - g["f"]).max() <= 1e-12 * fmax
    t = o.thermo()
    for k, key in enumerate(("temp", "pe", "ke", "press")):
        assert abs(t[key] - g["thermo0"][k]) <= 1e-11 * abs(g["thermo0"][k])
    series = []
    for _ in range(10):  # each reference `run 10` starts with Verlet::setup (rebuild + force evaluation)
        o.setup()
        series.append(o.run(10, s["dt"], 0)[-1])
    for row, t in zip(g["thermo_series"], series):
        for k, key in enumerate(("temp", "pe", "ke", "press")):
            assert abs(t[key] - row[k]) <= 1e-9 * abs(row[k]), (key, t[key], row[k])
    o.close()


@pytest.mark.parametrize("name,kind", CASES)
def test_full_list(pkg, oracle_mod, name, kind):
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    o, s = build_oracle(oracle_mod, pkg, g, kind, False)
    o.neigh_build()
    a = o.atoms()
    nn, flat = o.neigh_tags()
    got = Cm.full_sets(a["tag"], nn, flat)
    want = np.sort(g["full_i"].astype(np.int64) * (1 << 32) + g["full_j"].astype(np.int64))
    assert np.array_equal(got, want)
    o.close()


def ljmix_case(pkg, mix):
    """tests/golden/ljmix.npz (make_golden.py ljmix): two types, per-type cutoffs 2.5 / 2.0, the 1-2 pair mixed by
    PairLJCut::init_one (src/pair_lj_cut.cpp:501-556) -> epsilon, sigma AND cutoff from Pair::mix_*"""
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, "ljmix.npz"))
    T = B.lj_cut_tables(2, {(1, 1): (1.0, 1.0, 2.5), (2, 2): (0.5, 0.9, 2.0)}, 2.5, mix=mix)
    scale = (4.0 / 0.8442) ** (1.0 / 3.0)
    nat = len(g[mix + "_type"])
    s = dict(boxlo=[0.0] * 3, boxhi=[4 * scale] * 3, x=g[mix + "_x"], v=g[mix + "_v"], type=g[mix + "_type"],
             tag=np.arange(1, nat + 1, dtype=np.int32), mass=np.array([0.0, 1.0, 1.5]), units=B.UNITS["lj"],
             dt=0.005, skin=0.3)
    cn, cmax = B.cutneighsq_table(2, T["cutsq"], 0.3)
    return g, T, s, cn, cmax


@pytest.mark.parametrize("mix", ["geometric", "arithmetic"])
def test_lj_mixed_cutoffs(pkg, oracle_mod, mix):
    g, T, s, cn, cmax = ljmix_case(pkg, mix)
    c12 = np.sqrt(2.5 * 2.0) if mix == "geometric" else 2.25
    assert T["cutsq"][1, 2] == pytest.approx(c12 * c12, rel=1e-15)
    o = oracle_mod.Oracle(s["boxlo"], s["boxhi"], s["x"], s["v"], s["type"], s["tag"], s["mass"], s["units"])
    o.pair_lj(T)
    o.neigh_config(cn, cmax, 0.3, 20, 0, 0, True)
    o.setup()
    a = o.atoms()
    nn, flat = o.neigh_tags()
    want = np.sort(g[mix + "_half_lo"].astype(np.int64) * (1 << 32) + g[mix + "_half_hi"].astype(np.int64))
    assert np.array_equal(Cm.pairs_from_lists(a["tag"], nn, flat), want)
    f = a["f"][np.argsort(a["tag"])]
    assert np.abs(f - g[mix + "_f"]).max() <= 1e-12 * np.abs(g[mix + "_f"]).max()
    t = o.thermo()
    for k, key in enumerate(("temp", "pe", "ke", "press")):
        assert abs(t[key] - g[mix + "_thermo0"][k]) <= 1e-11 * abs(g[mix + "_thermo0"][k])
    t = o.run(20, 0.005, 0)[-1]
    for k, key in enumerate(("temp", "pe", "ke", "press")):
        assert abs(t[key] - g[mix + "_thermo20"][k]) <= 1e-9 * abs(g[mix + "_thermo20"][k])
    o.close()
