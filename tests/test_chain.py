"""bench/in.chain (FENE bead-spring melt, 32 000 atoms, `atom_style bond`, `special_bonds fene`,
lj/cut 1.12 shift yes + bond fene) WITHOUT `fix langevin` -- SURVEY.md 8(f) rank 2.
Golden: tests/golden/chain.npz from the unmodified reference (tests/golden/make_golden.py chain):
digests of the half pair list (1-2 pairs excluded by find_special), forces of every 8th tag,
thermo with evdwl / emol at step 0 and over 100 steps.  The step-0 numbers are also the ones the
reference's log prints (bench/log.6Oct16.chain.fixed.icc.1; langevin does not act before step 1).
data.chain itself is shipped beside the reference build (oracle/_ref/data/, not in the repo)."""
import os

import numpy as np
import pytest

import common as Cm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "oracle", "_ref", "data", "data.chain")
KEYS = ("temp", "pe", "ke", "press", "evdwl", "emol")
F_TOL, T_TOL = 1e-10, 1e-8


def golden():
    return np.load(os.path.join(Cm.GOLDEN, "chain.npz"))


def pair_hash(lo, hi):
    k = lo.astype(np.uint64) * np.uint64(1 << 32) + hi.astype(np.uint64)
    m = k * np.uint64(0x9E3779B97F4A7C15)
    m ^= m >> np.uint64(29)
    m *= np.uint64(0xBF58476D1CE4E5B9)
    return np.array([np.bitwise_xor.reduce(m), m.sum(dtype=np.uint64)], dtype=np.uint64)


def check_pairs(g, tags_i, numneigh, flat):
    i = np.repeat(tags_i.astype(np.int64), numneigh)
    j = flat.astype(np.int64)
    lo, hi = np.minimum(i, j), np.maximum(i, j)
    assert len(lo) == int(g["npairs"])
    cnt = np.zeros(int(g["natoms"]) + 1, np.int64)
    np.add.at(cnt, lo, 1); np.add.at(cnt, hi, 1)
    assert np.array_equal(cnt.astype(np.uint8), g["pairs_per_tag"])
    assert np.array_equal(pair_hash(lo, hi), g["pair_digest"])      # bit-exact pair set


def check_forces(g, tag, f):
    fo = np.empty_like(f); fo[tag - 1] = f
    fmax = np.abs(g["f_every8"]).max()
    assert np.abs(fo[::8] - g["f_every8"]).max() <= F_TOL * fmax
    assert np.allclose(np.abs(fo).sum(axis=0), g["f_abs_sum"], rtol=1e-11)
    assert abs((fo * fo).sum() - float(g["f_sq_sum"])) <= 1e-11 * float(g["f_sq_sum"])


def thermo_vec(t):
    return np.array([t["temp"], t["pe"], t["ke"], t["press"], t["pe"] - t["emol"], t["emol"]])


def check_thermo(got, want, tol=T_TOL):
    for k, key in enumerate(KEYS):
        assert abs(got[k] - want[k]) <= tol * abs(want[k]), (key, got[k], want[k])


def need_data():
    if not os.path.exists(DATA):
        pytest.skip("oracle/_ref/data/data.chain missing (oracle/build_ref.py copies it from the reference)")


def test_oracle_matches_reference(pkg, oracle_mod):
    need_data()
    g = golden()
    M, B = pkg.molecule_host, pkg.b200md
    d = M.read_data(DATA)
    tp = M.build_topology(d["natoms"], d["bonds"])
    o = oracle_mod.Oracle(d["boxlo"], d["boxhi"], d["x"], d["v"], d["type"], d["tag"], d["mass"], B.UNITS["lj"])
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 1.12)}, 1.12, shift=True)
    o.pair_lj(T)
    o.set_topology(tp)
    o.bond_fene([0.0, 30.0], [0.0, 1.5], [0.0, 1.0], [0.0, 1.0])
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.4)
    o.neigh_config(cn, cmax, 0.4, 1, 1, 1, True)
    o.setup()
    a = o.atoms()
    nn, flat = o.neigh_tags()
    check_pairs(g, a["tag"], nn, flat)
    check_forces(g, a["tag"], a["f"])
    check_thermo(thermo_vec(o.thermo()), g["thermo0"], 1e-11)
    for row in g["thermo_series"]:
        o.setup()
        check_thermo(thermo_vec(o.run(10, 0.012, 0)[-1]), row)
    o.close()


def test_oracle_replicated_matches_reference(pkg, oracle_mod):
    """in.chain.scaled -var x 2 -var y 1 -var z 2: host-side `replicate` + the oracle against the
    reference's thermo (128 000 atoms)"""
    need_data()
    g = golden()
    M, B = pkg.molecule_host, pkg.b200md
    d = M.replicate(M.read_data(DATA), 2, 1, 2)
    assert d["natoms"] == 128000 and len(d["bonds"]) == 4 * 31680
    tp = M.build_topology(d["natoms"], d["bonds"])
    o = oracle_mod.Oracle(d["boxlo"], d["boxhi"], d["x"], d["v"], d["type"], d["tag"], d["mass"], B.UNITS["lj"])
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 1.12)}, 1.12, shift=True)
    o.pair_lj(T); o.set_topology(tp)
    o.bond_fene([0.0, 30.0], [0.0, 1.5], [0.0, 1.0], [0.0, 1.0])
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.4)
    o.neigh_config(cn, cmax, 0.4, 1, 1, 1, True)
    o.setup()
    check_thermo(thermo_vec(o.thermo()), g["rep212_thermo0"], 1e-10)
    o.close()


def test_special_tables_match_reference_counts(pkg):
    """Ave special neighs/atom = 1.98 and 31680 bonds in the reference's log for data.chain"""
    need_data()
    M = pkg.molecule_host
    d = M.read_data(DATA)
    tp = M.build_topology(d["natoms"], d["bonds"])
    assert d["natoms"] == 32000 and len(d["bonds"]) == 31680
    assert tp["special_offset"][-1] == 2 * 31680 and abs(tp["special_offset"][-1] / 32000 - 1.98) < 1e-12
    assert tp["bond_offset"][-1] == 31680 and np.all(tp["special_level"] == 1)
    # 1-3 / 1-4 levels appear only when their weights differ from 1 (Special::build early return)
    tp2 = M.build_topology(d["natoms"], d["bonds"], special_lj=(0, 0, 0, 1), special_coul=(0, 0, 0, 1))
    assert set(int(v) for v in np.unique(tp2["special_level"])) == {1, 2, 3}
    chain = d["bonds"][d["bonds"][:, 1] == 2]
    assert len(chain) == 1 and chain[0, 2] == 3


@pytest.mark.gpu
@pytest.mark.parametrize("lanes,half", [(2, True), (4, True), (2, False)])
def test_cuda_matches_reference(pkg, lanes, half):
    need_data()
    g = golden()
    B = pkg.b200md
    ctx = B.Context(0)
    ctx.set_lanes_per_atom(lanes)
    cfg = pkg.inputs.in_chain(ctx, DATA, list_kind=B.LIST_HALF_NEWTON if half else B.LIST_FULL)
    ctx.setup()
    a = ctx.get_atoms(("tag", "f"))
    if half:
        nn, flat = ctx.neigh_export(by_tag=True)
        check_pairs(g, a["tag"], nn, flat)
    check_forces(g, a["tag"], a["f"])
    t0 = thermo_vec(ctx.thermo())
    check_thermo(t0, g["thermo0"], 1e-10)
    # the reference's own log at step 0: Temp E_pair E_mol TotEng Press (per atom)
    n = 32000.0
    log0 = np.array([0.97029772, 0.44484087, 20.494523, 22.394765, 4.6721833])
    mine = np.array([t0[0], t0[4] / n, t0[5] / n, (t0[1] + t0[2]) / n, t0[3]])
    assert np.allclose(mine, log0, rtol=6e-8)
    for row in g["thermo_series"]:
        ctx.setup()
        check_thermo(thermo_vec(ctx.run(10, cfg["dt"], 0)[-1]), row)
    ctx.close()


@pytest.mark.gpu
def test_cuda_replicated_matches_reference(pkg):
    need_data()
    g = golden()
    B = pkg.b200md
    ctx = B.Context(0)
    cfg = pkg.inputs.in_chain(ctx, DATA, 2, 1, 2)
    assert cfg["natoms"] == 128000
    ctx.setup()
    check_thermo(thermo_vec(ctx.thermo()), g["rep212_thermo0"], 1e-10)
    for row in g["rep212_series"]:
        ctx.setup()
        check_thermo(thermo_vec(ctx.run(10, cfg["dt"], 0)[-1]), row)
    ctx.close()


@pytest.mark.gpu
def test_cuda_chain_conserves_energy(pkg):
    """NVE property at the reference's time step: total energy drift over 200 steps < 2e-4 relative"""
    need_data()
    B = pkg.b200md
    ctx = B.Context(0)
    cfg = pkg.inputs.in_chain(ctx, DATA)
    ctx.setup()
    t0 = ctx.thermo()
    t1 = ctx.run(200, cfg["dt"], 0)[-1]
    e0, e1 = t0["pe"] + t0["ke"], t1["pe"] + t1["ke"]
    assert abs(e1 - e0) <= 2e-4 * abs(e0), (e0, e1)
    st = ctx.neigh_stats()
    assert st["nbuilds"] >= 10
    ctx.close()
