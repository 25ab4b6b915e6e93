"""fix langevin on the device path (bench/in.chain as the reference ships
it).  FixLangevin::post_force (src/fix_langevin.cpp:576-713) draws three RanMars::uniform()
numbers per atom in LOCAL INDEX order (src/random_mars.cpp:78-95), so parity needs (i) the
reference's local order — Atom::sort at Verlet::setup (src/atom.cpp:2246-2349, src/verlet.cpp:123)
— and (ii) the exact Marsaglia stream.  The oracle restates both serially; the CUDA path
generates the same stream in parallel by polynomial jump-ahead of the lagged-Fibonacci recurrence.
Golden: tests/golden/langevin.npz from the unmodified reference (tests/golden/make_langevin.py)
and the thermo lines of bench/log.6Oct16.chain.fixed.icc.1.
Tolerances: forces after set-up 1e-10 relative to max |f| (single evaluation); thermo 1e-8
relative over 100 steps; x, v after 40 steps 1e-9 relative to the box / max |v|."""
import os

import numpy as np
import pytest

import common as Cm

F_TOL, T_TOL = 1e-10, 1e-8
KEYS = ("temp", "pe", "ke", "press", "evdwl", "emol")
# bench/log.6Oct16.chain.fixed.icc.1: Step Temp E_pair E_mol TotEng Press at step 100, per atom
LOG_STEP100 = dict(temp=0.9729966, e_pair=0.4361122, e_mol=20.507698, press=4.6548819)


def golden():
    return np.load(os.path.join(Cm.GOLDEN, "langevin.npz"))


def tvec(t):
    return np.array([t["temp"], t["pe"], t["ke"], t["press"], t["pe"] - t["emol"], t["emol"]])


def close(got, want, tol, n=6):
    for k in range(n):
        assert abs(got[k] - want[k]) <= tol * abs(want[k]), (KEYS[k], got[k], want[k])


def lj_inputs(pkg, G):
    """the lj_n5 state in the reference's local order (the order its RanMars draws follow)"""
    order = G["lj_local_tags"].astype(np.int64) - 1
    s = Cm.lj_system(pkg, 1)
    scale = (4.0 / 0.8442) ** (1.0 / 3.0)
    s["boxlo"], s["boxhi"] = [0.0] * 3, [5 * scale] * 3
    s.update(x=G["lj_x0"][order], v=G["lj_v0"][order], tag=(order + 1).astype(np.int32),
             type=np.ones(len(order), np.int32))
    return s


def check_lj_final(G, s, tag, x, v, f, image, thermo):
    o = np.argsort(tag)
    prd = np.array(s["boxhi"]) - np.array(s["boxlo"])
    d = x[o] - G["lj_x"]
    d -= prd * np.round(d / prd)           # the two runs start from different image flags
    assert np.abs(d).max() <= 1e-9 * prd[0]
    assert np.abs(v[o] - G["lj_v"]).max() <= 1e-9 * np.abs(G["lj_v"]).max()
    assert np.abs(f[o] - G["lj_f"]).max() <= 1e-8 * np.abs(G["lj_f"]).max()
    close(tvec(thermo), G["lj_thermo"], T_TOL, 4)


# ---- the stream itself ------------------------------------------------------------------------
def test_ranmars_integer_form(oracle_mod):
    """RanMars is exact 24-bit integer arithmetic: L[n+97] = L[n] - L[n+64] (mod 2^24) and
    c[n] = (362436 - (n+1)*7654321) mod 16777213 — the closed form the CUDA generator is built on."""
    seed, n = 904297, 5000
    u = oracle_mod.ranmars_stream(seed, n)
    ui = np.round(u * (1 << 24)).astype(np.int64)
    assert np.array_equal(ui / float(1 << 24), u)
    c = (362436 - ((np.arange(n, dtype=np.int64) + 2) * 7654321)) % 16777213   # draw k is call k+1 (the constructor drew once)
    w = (ui + c) % (1 << 24)                                                     # the lagged-Fibonacci values
    assert np.array_equal(w[97:], (w[:-97] - w[64:-33]) % (1 << 24))


def test_ranmars_jump_ahead(pkg, oracle_mod):
    """host tables of the device generator (x^k mod x^97 + x^64 - 1 over Z/2^24) reproduce the
    serial stream at arbitrary offsets"""
    B = pkg.b200md
    for seed in (48271, 904297, 1, 900000000):
        u = oracle_mod.ranmars_stream(seed, 40000)
        st = B.ranmars_state(seed)               # position 1: the constructor's draw is consumed
        for k in (0, 1, 96, 97, 2048, 2049, 30000, 39800):
            got = B.ranmars_draws(B.ranmars_jump(st, k), k + 1, 200)
            assert np.array_equal(got, u[k:k + 200]), (seed, k)
    # a jump of 2^40 draws equals two jumps of 2^39 (the tables the device composes)
    st = B.ranmars_state(12345)
    assert np.array_equal(B.ranmars_jump(st, 1 << 40), B.ranmars_jump(B.ranmars_jump(st, 1 << 39), 1 << 39))


# ---- oracle vs the unmodified reference --------------------------------------------------------
def chain_oracle(pkg, oracle_mod):
    import test_chain as Tc
    Tc.need_data()
    M, B = pkg.molecule_host, pkg.b200md
    d = M.read_data(Tc.DATA)
    tp = M.build_topology(d["natoms"], d["bonds"])
    o = oracle_mod.Oracle(d["boxlo"], d["boxhi"], d["x"], d["v"], d["type"], d["tag"], d["mass"], B.UNITS["lj"])
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 1.12)}, 1.12, shift=True)
    o.pair_lj(T); o.set_topology(tp)
    o.bond_fene([0.0, 30.0], [0.0, 1.5], [0.0, 1.0], [0.0, 1.0])
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.4)
    o.neigh_config(cn, cmax, 0.4, 1, 1, 1, True)
    o.atom_sort()
    o.fix_langevin(1.0, 1.0, 10.0, 904297, 0.012)
    return o


def check_chain(G, after_setup, series, final):
    tag, f = after_setup
    fo = np.empty_like(f); fo[tag - 1] = f
    fmax = np.abs(G["chain_f_every8"]).max()
    assert np.abs(fo[::8] - G["chain_f_every8"]).max() <= F_TOL * fmax
    assert np.allclose(np.abs(fo).sum(axis=0), G["chain_f_abs_sum"], rtol=1e-11)
    for row, t in zip(G["chain_series"], series):
        close(tvec(t), row, T_TOL)
    t, n = series[-1], 32000
    for key, got in (("temp", t["temp"]), ("e_pair", (t["pe"] - t["emol"]) / n), ("e_mol", t["emol"] / n), ("press", t["press"])):
        assert abs(got - LOG_STEP100[key]) <= 6e-8 * abs(LOG_STEP100[key]), key   # the log prints 8 digits
    tag, x, v = final
    xo = np.empty_like(x); xo[tag - 1] = x
    vo = np.empty_like(v); vo[tag - 1] = v
    assert np.abs(vo[::8] - G["chain_v_every8"]).max() <= 1e-8 * np.abs(G["chain_v_every8"]).max()
    assert np.abs(xo[::8] - G["chain_x_every8"]).max() <= 1e-8 * np.abs(G["chain_x_every8"]).max()


def test_oracle_chain_langevin(pkg, oracle_mod):
    G = golden()
    o = chain_oracle(pkg, oracle_mod)
    o.setup()
    a = o.atoms()
    close(tvec(o.thermo()), G["chain_thermo0"], 1e-11)
    after = (a["tag"].copy(), a["f"].copy())
    series = [o.run(10, 0.012, 0)[-1] for _ in range(10)]
    a = o.atoms()
    check_chain(G, after, series, (a["tag"], a["x"], a["v"]))
    o.close()


def test_oracle_chain_replicated_langevin(pkg, oracle_mod):
    """in.chain.scaled 2x1x2: `replicate` leaves the atoms in creation order (src/replicate.cpp:431-640),
    Atom::sort is stable on top of it, and the draws follow"""
    import test_chain as Tc
    Tc.need_data()
    G = golden()
    M, B = pkg.molecule_host, pkg.b200md
    d = M.replicate(M.read_data(Tc.DATA), 2, 1, 2)
    lo = d["local_order"]
    tp = M.build_topology(d["natoms"], d["bonds"])
    o = oracle_mod.Oracle(d["boxlo"], d["boxhi"], d["x"][lo], d["v"][lo], d["type"][lo], d["tag"][lo], d["mass"], B.UNITS["lj"])
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 1.12)}, 1.12, shift=True)
    o.pair_lj(T); o.set_topology(tp)
    o.bond_fene([0.0, 30.0], [0.0, 1.5], [0.0, 1.0], [0.0, 1.0])
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.4)
    o.neigh_config(cn, cmax, 0.4, 1, 1, 1, True)
    o.atom_sort()
    o.fix_langevin(1.0, 1.0, 10.0, 904297, 0.012)
    o.setup()
    close(tvec(o.run(20, 0.012, 0)[-1]), G["rep212_thermo20"], 1e-9)
    o.close()


def test_oracle_lj_langevin_ramp(pkg, oracle_mod):
    G = golden()
    B = pkg.b200md
    s = lj_inputs(pkg, G)
    o = oracle_mod.Oracle(s["boxlo"], s["boxhi"], s["x"], s["v"], s["type"], s["tag"], s["mass"], s["units"])
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    o.pair_lj(T)
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.3)
    o.neigh_config(cn, cmax, 0.3, 20, 0, 0, True)
    o.atom_sort()
    o.fix_langevin(1.0, 2.5, 0.5, 48271, 0.005, ratio=[0.0, 1.5])
    o.setup()
    t = o.run(40, 0.005, 0)[-1]
    a = o.atoms()
    check_lj_final(G, s, a["tag"], a["x"], a["v"], a["f"], a["image"], t)
    o.close()


def ljsort_inputs(pkg, G):
    order = G["ljsort_local_tags"].astype(np.int64) - 1
    s = Cm.lj_system(pkg, 1)
    scale = (4.0 / 0.8442) ** (1.0 / 3.0)
    s["boxlo"], s["boxhi"] = [0.0] * 3, [5 * scale] * 3
    s.update(x=G["ljsort_x0"][order], v=G["ljsort_v0"][order], tag=(order + 1).astype(np.int32),
             type=np.ones(len(order), np.int32))
    return s


def check_ljsort(G, s, tag, x, v, thermo, xtol, ttol):
    o = np.argsort(tag)
    prd = np.array(s["boxhi"]) - np.array(s["boxlo"])
    d = x[o] - G["ljsort_x"]
    d -= prd * np.round(d / prd)
    assert np.abs(d).max() <= xtol * prd[0]
    assert np.abs(v[o] - G["ljsort_v"]).max() <= xtol * np.abs(G["ljsort_v"]).max()
    close(tvec(thermo), G["ljsort_thermo"], ttol, 4)


def test_oracle_langevin_across_resort(pkg, oracle_mod):
    """ONE run of 1100 steps: at the first reneighboring at or after step 1000 the reference re-sorts
    its atoms (Atom::sort, src/verlet.cpp:285) and the draws follow the new order.  A stream handed out
    in the old order shows up as O(1e-2) in the thermo; the restatement stays at rounding level."""
    G = golden()
    B = pkg.b200md
    s = ljsort_inputs(pkg, G)
    o = oracle_mod.Oracle(s["boxlo"], s["boxhi"], s["x"], s["v"], s["type"], s["tag"], s["mass"], s["units"])
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    o.pair_lj(T)
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.3)
    o.neigh_config(cn, cmax, 0.3, 20, 0, 0, True)
    o.atom_sort()
    o.fix_langevin(1.0, 1.0, 1.0, 7771, 0.005)
    o.setup()
    t = o.run(1100, 0.005, 0)[-1]
    a = o.atoms()
    check_ljsort(G, s, a["tag"], a["x"], a["v"], t, 1e-7, 1e-7)
    assert np.array_equal(a["tag"], G["ljsort_local_tags_end"])       # the reference's local order after its re-sort
    o.close()


def test_host_atom_sort_matches_oracle(pkg, oracle_mod):
    """the host mirror of Atom::sort (what the Python driver applies before the upload) gives the
    oracle's order"""
    G = golden()
    B = pkg.b200md
    s = lj_inputs(pkg, G)
    rng = np.random.default_rng(5)
    p = rng.permutation(len(s["tag"]))
    o = oracle_mod.Oracle(s["boxlo"], s["boxhi"], s["x"][p], s["v"][p], s["type"][p], s["tag"][p], s["mass"], s["units"])
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.3)
    o.neigh_config(cn, cmax, 0.3, 20, 0, 0, True)
    o.atom_sort()
    perm = pkg.inputs.atom_sort_order(s["x"][p], s["boxlo"], s["boxhi"], cmax)
    assert np.array_equal(o.atoms()["tag"], s["tag"][p][perm])
    o.close()


# ---- CUDA path through the C ABI ---------------------------------------------------------------
@pytest.mark.gpu
def test_cuda_ranmars_stream(pkg, oracle_mod):
    B = pkg.b200md
    ctx = B.Context(0)
    for seed, n in ((904297, 96000), (48271, 5), (1, 2048), (900000000, 300001)):
        u = oracle_mod.ranmars_stream(seed, 2 * n)
        got = ctx.ranmars_stream(seed, n, 2)           # two consecutive fills of n draws each
        assert np.array_equal(got, u), (seed, n)
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("half", [True, False])
def test_cuda_chain_langevin(pkg, half):
    import test_chain as Tc
    Tc.need_data()
    G = golden()
    B = pkg.b200md
    ctx = B.Context(0)
    pkg.inputs.in_chain(ctx, Tc.DATA, list_kind=B.LIST_HALF_NEWTON if half else B.LIST_FULL, langevin=True)
    ctx.setup()
    a = ctx.get_atoms()
    close(tvec(ctx.thermo()), G["chain_thermo0"], 1e-10)
    after = (a["tag"].copy(), a["f"].copy())
    series = [ctx.run(10, 0.012, 0)[-1] for _ in range(10)]
    a = ctx.get_atoms()
    check_chain(G, after, series, (a["tag"], a["x"], a["v"]))
    ctx.close()


@pytest.mark.gpu
def test_cuda_chain_replicated_langevin(pkg):
    import test_chain as Tc
    Tc.need_data()
    G = golden()
    B = pkg.b200md
    ctx = B.Context(0)
    pkg.inputs.in_chain(ctx, Tc.DATA, 2, 1, 2, langevin=True)
    ctx.setup()
    close(tvec(ctx.run(20, 0.012, 0)[-1]), G["rep212_thermo20"], T_TOL)
    ctx.close()


@pytest.mark.gpu
@pytest.mark.parametrize("half", [True, False])
def test_cuda_lj_langevin_ramp(pkg, half):
    G = golden()
    B = pkg.b200md
    s = lj_inputs(pkg, G)
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], 0.3)
    perm = pkg.inputs.atom_sort_order(s["x"], s["boxlo"], s["boxhi"], cmax)
    ctx = B.Context(0)
    ctx.set_units("lj"); ctx.set_box(s["boxlo"], s["boxhi"])
    ctx.set_atoms(s["x"][perm], s["v"][perm], s["type"][perm], s["tag"][perm], s["mass"])
    ctx.pair_lj_cut_config(*[T[k] for k in ("cutsq", "lj1", "lj2", "lj3", "lj4", "offset")])
    ctx.neigh_config(cn, cmax, 0.3, every=20, delay=0, check=0, list_kind=B.LIST_HALF_NEWTON if half else B.LIST_FULL)
    ctx.fix_langevin(1.0, 2.5, 0.5, 48271, 0.005, ratio=[0.0, 1.5])
    ctx.setup()
    t = ctx.run(40, 0.005, 0)[-1]
    a = ctx.get_atoms(("x", "v", "f", "tag", "image"))
    check_lj_final(G, s, a["tag"], a["x"], a["v"], a["f"], a["image"], t)
    ctx.close()


@pytest.mark.gpu
def test_cuda_langevin_refusals(pkg):
    B = pkg.b200md
    s = Cm.lj_system(pkg, 4)
    ctx = B.Context(0)
    ctx.set_units("lj"); ctx.set_box(s["boxlo"], s["boxhi"])
    with pytest.raises(RuntimeError, match="before|atoms"):
        ctx.fix_langevin(1.0, 1.0, 1.0, 1, 0.005)
    ctx.set_atoms(s["x"], s["v"], s["type"], s["tag"], s["mass"])
    with pytest.raises(RuntimeError, match="period"):
        ctx.fix_langevin(1.0, 1.0, 0.0, 1, 0.005)
    with pytest.raises(RuntimeError, match="seed"):
        ctx.fix_langevin(1.0, 1.0, 1.0, 0, 0.005)
    ctx.close()


# ---- properties of the host-side stream helpers (no GPU) -----------------------------------------
def test_ranmars_jump_properties(pkg, oracle_mod):
    """random seeds and offsets: jump(k) then draw == the serial stream at offset k; jumps compose"""
    B = pkg.b200md
    rng = np.random.default_rng(20231105)
    for _ in range(12):
        seed = int(rng.integers(1, 900000001))
        k1, k2 = int(rng.integers(0, 15000)), int(rng.integers(0, 15000))
        u = oracle_mod.ranmars_stream(seed, k1 + k2 + 64)
        st = B.ranmars_state(seed)
        a = B.ranmars_jump(st, k1)
        assert np.array_equal(B.ranmars_draws(a, 1 + k1, 64), u[k1:k1 + 64])
        b = B.ranmars_jump(a, k2)
        assert np.array_equal(b, B.ranmars_jump(st, k1 + k2))
        assert np.array_equal(B.ranmars_draws(b, 1 + k1 + k2, 64), u[k1 + k2:k1 + k2 + 64])
    with pytest.raises(B.B200MDError, match="Invalid seed"):
        B.ranmars_state(0)
    with pytest.raises(B.B200MDError, match="Invalid seed"):
        B.ranmars_state(900000001)


def test_atom_sort_order_properties(pkg):
    """a permutation, stable inside a bin, bins ascending (x fastest), atoms outside the box clamped"""
    B = pkg.b200md
    rng = np.random.default_rng(7)
    lo, hi = np.array([0.0, -1.0, 2.0]), np.array([9.0, 4.0, 7.5])
    x = lo + (hi - lo) * rng.random((5000, 3))
    x[:50] += 0.3 * (hi - lo) * rng.standard_normal((50, 3))        # some atoms outside the box
    cmax = 2.8
    perm = B.atom_sort_order(x, lo, hi, cmax)
    assert np.array_equal(np.sort(perm), np.arange(len(x)))
    nb = np.maximum(((hi - lo) / (0.5 * cmax)).astype(int), 1)
    c = np.clip(((x - lo) * (nb / (hi - lo))).astype(int), 0, nb - 1)
    bins = (c[:, 2] * nb[1] + c[:, 1]) * nb[0] + c[:, 0]
    sb = bins[perm]
    assert np.all(np.diff(sb) >= 0)
    same = np.diff(sb) == 0
    assert np.all(np.diff(perm)[same] > 0)                           # stable
    assert np.array_equal(B.atom_sort_order(x, lo, hi, 0.0), np.arange(len(x)))   # sorting disabled
