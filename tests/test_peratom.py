"""Per-atom energy / virial: the eflag_atom / vflag_atom branches of
Pair::ev_tally (src/pair.cpp:1036-1040, 1078-1097), PairEAM's eatom += phi
(src/MANYBODY/pair_eam.cpp:245) and Bond::ev_tally (src/bond.cpp:186-251), summed and
reverse-communicated the way `compute pe/atom` / `compute stress/atom` consume them
(src/compute_pe_atom.cpp:92-170, src/compute_stress_atom.cpp:140-340).
Golden: tests/golden/peratom.npz from the unmodified reference (tests/golden/make_peratom.py) at
the states of lj_n5 / eam_n4 / eamalloy_n4 and data.chain (every 8th tag).
Tolerance: 1e-10 relative to the largest per-atom magnitude (single evaluation, like forces)."""
import os

import numpy as np
import pytest

import common as Cm
from test_oracle_golden import CASES, build_oracle

TOL = 1e-10


def golden():
    return np.load(os.path.join(Cm.GOLDEN, "peratom.npz"))


def by_tag(tag, a):
    out = np.empty_like(a)
    out[tag - 1] = a
    return out


def check(P, name, tag, eatom, vatom, v, mass_of_atom, units, sel=slice(None)):
    """eatom / vatom in local order -> tag order -> the three reference computes"""
    e = by_tag(tag, eatom)
    va = by_tag(tag, vatom)
    want = P[name + "_pe"]
    assert np.abs(e[sel] - want).max() <= TOL * np.abs(want).max()
    assert abs(e.sum() - float(P[name + "_pe_sum"])) <= 1e-11 * abs(float(P[name + "_pe_sum"]))
    assert abs(e.sum() - float(P[name + "_thermo_pe"])) <= 1e-11 * abs(float(P[name + "_thermo_pe"]))
    # compute stress/atom NULL virial = -nktv2p * vatom
    sv = -units["nktv2p"] * va
    want = P[name + "_stress_virial"]
    assert np.abs(sv[sel] - want).max() <= TOL * np.abs(want).max()
    assert np.allclose(sv.sum(axis=0), P[name + "_stress_virial_sum"], rtol=1e-10, atol=1e-10 * np.abs(want).max())
    # compute stress/atom NULL adds the kinetic term mvv2e*m*v_a*v_b (compute_stress_atom.cpp:263-290)
    vt = by_tag(tag, v)
    m = by_tag(tag, mass_of_atom)[:, None] * units["mvv2e"]
    ke = np.stack([vt[:, 0] * vt[:, 0], vt[:, 1] * vt[:, 1], vt[:, 2] * vt[:, 2],
                   vt[:, 0] * vt[:, 1], vt[:, 0] * vt[:, 2], vt[:, 1] * vt[:, 2]], axis=1) * m
    s = -units["nktv2p"] * (va + ke)
    want = P[name + "_stress"]
    assert np.abs(s[sel] - want).max() <= TOL * np.abs(want).max()
    # the pressure identity the reference documents for stress/atom: P = -sum(sxx+syy+szz)/(3 V)
    press = -(s[:, 0] + s[:, 1] + s[:, 2]).sum() / (3.0 * float(P[name + "_vol"]))
    assert abs(press - float(P[name + "_thermo_press"])) <= 1e-9 * abs(float(P[name + "_thermo_press"]))


@pytest.mark.parametrize("name,kind", CASES)
def test_oracle_peratom(pkg, oracle_mod, name, kind):
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    P = golden()
    o, s = build_oracle(oracle_mod, pkg, g, kind, True)
    o.setup()
    eatom, vatom = o.peratom()
    a = o.atoms()
    check(P, name, a["tag"], eatom, vatom, a["v"], s["mass"][a["type"]], s["units"])
    o.close()


def test_oracle_peratom_chain(pkg, oracle_mod):
    import test_chain as Tc
    Tc.need_data()
    P = golden()
    M, B = pkg.molecule_host, pkg.b200md
    d = M.read_data(Tc.DATA)
    tp = M.build_topology(d["natoms"], d["bonds"])
    o = oracle_mod.Oracle(d["boxlo"], d["boxhi"], d["x"], d["v"], d["type"], d["tag"], d["mass"], B.UNITS["lj"])
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 1.12)}, 1.12, shift=True)
    o.pair_lj(T); o.set_topology(tp)
    o.bond_fene([0.0, 30.0], [0.0, 1.5], [0.0, 1.0], [0.0, 1.0])
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.4)
    o.neigh_config(cn, cmax, 0.4, 1, 1, 1, True)
    o.setup()
    eatom, vatom = o.peratom()
    a = o.atoms()
    check(P, "chain", a["tag"], eatom, vatom, a["v"], np.asarray(d["mass"])[a["type"]], B.UNITS["lj"], slice(None, None, 8))
    o.close()


# ---- CUDA path through the C ABI -------------------------------------------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("half", [True, False])
@pytest.mark.parametrize("name,kind", CASES)
def test_cuda_peratom(pkg, oracle_mod, name, kind, half):
    from test_gpu_parity import make_ctx
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    P = golden()
    ctx, s = make_ctx(pkg, oracle_mod, g, kind, B.LIST_HALF_NEWTON if half else B.LIST_FULL, 2)
    ctx.setup()
    eatom, vatom = ctx.peratom()
    a = ctx.get_atoms()
    check(P, name, a["tag"], eatom, vatom, a["v"], s["mass"][a["type"]], s["units"])
    # energy-only and virial-only requests give the same numbers; the forces are untouched
    e2, _ = ctx.peratom(eflag_atom=1, vflag_atom=0)
    _, v2 = ctx.peratom(eflag_atom=0, vflag_atom=1)
    assert np.array_equal(e2, eatom) or np.abs(e2 - eatom).max() <= 1e-13 * np.abs(eatom).max()
    assert np.abs(v2 - vatom).max() <= 1e-13 * np.abs(vatom).max()
    assert np.array_equal(ctx.get_atoms(("f",))["f"], a["f"])
    # ... and after 10 more steps the per-atom sums still equal the global tallies
    t = ctx.run(10, s["dt"], 0)[-1]
    eatom, vatom = ctx.peratom()
    assert abs(eatom.sum() - t["pe"]) <= 1e-10 * abs(t["pe"])
    vir = np.array(t["virial"])
    assert np.allclose(vatom.sum(axis=0), vir, rtol=1e-9, atol=1e-9 * np.abs(vir).max())
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("half", [True, False])
def test_cuda_peratom_chain(pkg, half):
    import test_chain as Tc
    Tc.need_data()
    P = golden()
    B = pkg.b200md
    ctx = B.Context(0)
    cfg = pkg.inputs.in_chain(ctx, Tc.DATA, list_kind=B.LIST_HALF_NEWTON if half else B.LIST_FULL)
    ctx.setup()
    eatom, vatom = ctx.peratom()
    a = ctx.get_atoms()
    check(P, "chain", a["tag"], eatom, vatom, a["v"], np.ones(len(a["tag"])), B.UNITS["lj"], slice(None, None, 8))
    # pair-only and bond-only parts add up (compute pe/atom pair / bond keywords)
    ep, vp = ctx.peratom(which=B.PERATOM_PAIR)
    eb, vb = ctx.peratom(which=B.PERATOM_BOND)
    assert np.abs(ep + eb - eatom).max() <= 1e-12 * np.abs(eatom).max()
    assert np.abs(vp + vb - vatom).max() <= 1e-12 * np.abs(vatom).max()
    t = ctx.thermo()
    assert abs(eb.sum() - t["emol"]) <= 1e-10 * abs(t["emol"])
    ctx.close()
