"""GPU (-m gpu): the CUDA path through the C ABI (libb200md.so via ctypes) against
  (1) the golden vectors generated from the unmodified reference (tests/golden/*.npz),
  (2) the C oracle on seeded inputs at sizes it finishes in seconds,
  (3) the thermo the reference's own logs print for bench/in.lj and bench/in.eam (32k atoms),
  (4) size-independent properties at the bench's full 2M-atom size.
Tolerances (BASELINE.json north_star): neighbor lists bit-exact as sorted per-atom sets;
forces 1e-10 relative (to max |f|) on a single step; thermo 1e-8 relative over 100 steps.
"""
import os

import numpy as np
import pytest

import common as Cm
from test_oracle_golden import CASES, build_oracle, system_from_golden, tables_for

pytestmark = pytest.mark.gpu
F_TOL, T_TOL = 1e-10, 1e-8
KEYS = ("temp", "pe", "ke", "press")


def make_ctx(pkg, O, g, kind, list_kind, tpa=8):
    B = pkg.b200md
    T = tables_for(pkg, O, kind)
    s = system_from_golden(pkg, g, kind)
    if kind != "lj":
        s["mass"] = np.array([0.0, T["mass"][1]])
    ctx = B.Context(0)
    ctx.set_lanes_per_atom(tpa)
    ctx.set_units("lj" if kind == "lj" else "metal")
    ctx.set_box(s["boxlo"], s["boxhi"])
    ctx.set_atoms(s["x"], s["v"], s["type"], s["tag"], s["mass"], image=g["image"])
    if kind == "lj":
        ctx.pair_lj_cut_config(*[T[k] for k in ("cutsq", "lj1", "lj2", "lj3", "lj4", "offset")])
        cutsq = T["cutsq"]; sched = (20, 0, 0)
    else:
        ctx.pair_eam_config(T)
        cutsq = np.zeros((2, 2)); cutsq[1, 1] = T["cutforcesq"]; sched = (1, 5, 1)
    cn, cmax = B.cutneighsq_table(1, cutsq, s["skin"])
    ctx.neigh_config(cn, cmax, s["skin"], every=sched[0], delay=sched[1], check=sched[2], list_kind=list_kind)
    return ctx, s


@pytest.mark.parametrize("tpa", [1, 2, 4, 8, 16, 32])
@pytest.mark.parametrize("name,kind", CASES)
def test_golden_full_list_forces_thermo(pkg, oracle_mod, name, kind, tpa):
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    ctx, s = make_ctx(pkg, oracle_mod, g, kind, B.LIST_FULL, tpa)
    ctx.setup()
    a = ctx.get_atoms()
    nn, flat = ctx.neigh_export(by_tag=True)
    got = Cm.full_sets(a["tag"], nn, flat)
    want = np.sort(g["full_i"].astype(np.int64) * (1 << 32) + g["full_j"].astype(np.int64))
    assert np.array_equal(got, want)  # bit-exact per-atom sets
    order = np.argsort(a["tag"])
    fmax = np.abs(g["f"]).max()
    assert np.abs(a["f"][order] - g["f"]).max() <= F_TOL * fmax
    t = ctx.thermo()
    for k, key in enumerate(KEYS):
        assert abs(t[key] - g["thermo0"][k]) <= T_TOL * abs(g["thermo0"][k]), key
    for row in g["thermo_series"]:
        ctx.setup()
        t = ctx.run(10, s["dt"], 0)[-1]
        for k, key in enumerate(KEYS):
            assert abs(t[key] - row[k]) <= T_TOL * abs(row[k]), (key, t[key], row[k])
    ctx.close()


@pytest.mark.parametrize("tpa", [1, 2, 4, 16])
@pytest.mark.parametrize("name,kind", CASES)
def test_golden_half_list_forces_thermo(pkg, oracle_mod, name, kind, tpa):
    """half/bin/atomonly/newton (the reference's default list) for lj/cut, eam and eam/alloy:
    pair membership bit-exact, forces after reverse_comm, thermo series."""
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    ctx, s = make_ctx(pkg, oracle_mod, g, kind, B.LIST_HALF_NEWTON, tpa)
    ctx.setup()
    a = ctx.get_atoms()
    nn, flat = ctx.neigh_export(by_tag=True)
    got = Cm.pairs_from_lists(a["tag"], nn, flat)
    want = np.sort(g["half_lo"].astype(np.int64) * (1 << 32) + g["half_hi"].astype(np.int64))
    assert np.array_equal(got, want)
    order = np.argsort(a["tag"])
    assert np.abs(a["f"][order] - g["f"]).max() <= F_TOL * np.abs(g["f"]).max()
    t = ctx.thermo()
    for k, key in enumerate(KEYS):
        assert abs(t[key] - g["thermo0"][k]) <= T_TOL * abs(g["thermo0"][k]), key
    for row in g["thermo_series"]:
        ctx.setup()
        t = ctx.run(10, s["dt"], 0)[-1]
        for k, key in enumerate(KEYS):
            assert abs(t[key] - row[k]) <= T_TOL * abs(row[k]), (key, t[key], row[k])
    ctx.close()


def test_lattice_and_velocity_match_oracle_input(pkg, oracle_mod):
    """x, v after `velocity ... loop geom` + zero momentum + rescale: GPU state vs oracle state."""
    B = pkg.b200md
    s = Cm.lj_system(pkg, 8)
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    o = Cm.make_oracle(oracle_mod, pkg, s, "lj", T, 20, 0, 0)
    ctx = B.Context(0)
    pkg.inputs.in_lj(ctx, 1, 1, 1) if False else None
    ctx.set_units("lj"); ctx.set_box(s["boxlo"], s["boxhi"])
    ctx.set_atoms(s["x"], s["v"], s["type"], s["tag"], s["mass"])
    ctx.velocity_finalize(1.44, True)
    a, b = ctx.get_atoms(), o.atoms()
    assert np.array_equal(a["x"], b["x"])
    assert np.abs(a["v"] - b["v"]).max() <= 1e-13 * np.abs(b["v"]).max()
    ctx.close(); o.close()


@pytest.mark.parametrize("kind", ["lj", "eam"])
def test_32k_vs_oracle(pkg, oracle_mod, kind):
    """bench size (32 000 atoms): lists, single-step forces and 100-step thermo vs the oracle."""
    B = pkg.b200md
    ctx = B.Context(0)
    if kind == "lj":
        cfg = pkg.inputs.in_lj(ctx, 1, 1, 1)
        s = Cm.lj_system(pkg, 20)
        T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
        sched = (20, 0, 0)
    else:
        pot = os.path.join(oracle_mod.POT_DIR, "Cu_u3.eam")
        cfg = pkg.inputs.in_eam(ctx, 1, 1, 1, potential=pot)
        T = pkg.eam_host.tables_from_funcfl([pot])
        s = Cm.eam_system(pkg, 20, T["mass"][1])
        sched = (1, 5, 1)
    o = Cm.make_oracle(oracle_mod, pkg, s, kind, T, *sched, half=True)
    ofull = Cm.make_oracle(oracle_mod, pkg, s, kind, T, *sched, half=False)
    ctx.setup(); o.setup(); ofull.neigh_build()
    a, b = ctx.get_atoms(), o.atoms()
    nn, flat = ctx.neigh_export(True)
    nno, flo = ofull.neigh_tags()
    assert np.array_equal(Cm.full_sets(a["tag"], nn, flat), Cm.full_sets(ofull.atoms()["tag"], nno, flo))
    # a perfect lattice has zero net force: step 30 steps first so forces are O(10)
    ctx.run(30, cfg["dt"], 0); o.run(30, cfg["dt"], 0)
    a, b = ctx.get_atoms(), o.atoms()
    fa, fb = a["f"][np.argsort(a["tag"])], b["f"][np.argsort(b["tag"])]
    assert np.abs(fa - fb).max() <= 1e-9 * np.abs(fb).max()  # 30 steps of round-off growth
    ta = ctx.run(70, cfg["dt"], 0)[-1]; tb = o.run(70, cfg["dt"], 0)[-1]
    for key in KEYS:
        assert abs(ta[key] - tb[key]) <= T_TOL * abs(tb[key]), (key, ta[key], tb[key])
    assert ctx.neigh_stats()["nbuilds"] == o.nbuilds()
    ctx.close(); o.close(); ofull.close()


# thermo printed by the reference's own logs (bench/log.6Oct16.{lj,eam}.fixed.icc.1), 8 digits
LOG_LJ = {0: (1.44, -6.7733681, -4.6134356, -5.0197073), 100: (0.7574531, -5.7585055, -4.6223613, 0.20726105)}
LOG_EAM = {0: (1600.0, -113280.0, -106662.09, 18703.573), 50: (781.69049, -109873.35, -106640.13, 52273.088),
           100: (801.832, -109957.3, -106640.77, 51322.821)}


def _fmt8(v):
    return float("%.8g" % v)


@pytest.mark.parametrize("kind", ["lj", "eam"])
def test_bench_logs_reproduced(pkg, oracle_mod, kind):
    B = pkg.b200md
    ctx = B.Context(0)
    if kind == "lj":
        cfg = pkg.inputs.in_lj(ctx); log = LOG_LJ; per = 32000.0
    else:
        cfg = pkg.inputs.in_eam(ctx, potential=os.path.join(oracle_mod.POT_DIR, "Cu_u3.eam")); log = LOG_EAM; per = 1.0
    ctx.setup()
    recs = {0: ctx.thermo()}
    for t in ctx.run(100, cfg["dt"], 50):
        recs[t["step"]] = t
    for step, (temp, epair, etot, press) in log.items():
        t = recs[step]
        got = (t["temp"], t["pe"] / per, (t["pe"] + t["ke"]) / per, t["press"])
        for a, b in zip(got, (temp, epair, etot, press)):
            assert _fmt8(a) == pytest.approx(b, rel=2e-8), (step, got)
    ctx.close()


def test_full_size_properties(pkg):
    """2 048 000 atoms (the bench workload): list symmetry checksum, zero net force, energy
    conservation, rebuild count — properties that need no oracle."""
    B = pkg.b200md
    ctx = B.Context(0)
    cfg = pkg.inputs.in_lj(ctx, 4, 4, 4)
    ctx.setup()
    a = ctx.get_atoms(("tag",))
    nn, flat = ctx.neigh_export(True)
    i = np.repeat(a["tag"].astype(np.int64), nn); j = flat.astype(np.int64)
    # full list is symmetric: multiset of (i,j) equals multiset of (j,i) -> order-free checksums agree
    iu, ju, K = i.astype(np.uint64), j.astype(np.uint64), np.uint64(0x9E3779B97F4A7C15)
    h1 = np.bitwise_xor.reduce((iu * np.uint64(1000003) + ju) * K)  # uint64 wrap-around hash
    h2 = np.bitwise_xor.reduce((ju * np.uint64(1000003) + iu) * K)
    s1 = ((iu * np.uint64(1000003) + ju) * K).sum(); s2 = ((ju * np.uint64(1000003) + iu) * K).sum()
    assert h1 == h2 and s1 == s2
    assert 70.0 < len(j) / cfg["natoms"] < 80.0  # 78 on the perfect lattice, 75.1 in the melt
    t0 = ctx.thermo()
    t1 = ctx.run(100, cfg["dt"], 0)[-1]
    f = ctx.get_atoms(("f",))["f"]
    assert np.abs(f.sum(axis=0)).max() <= 1e-9 * np.abs(f).sum()
    e0, e1 = t0["pe"] + t0["ke"], t1["pe"] + t1["ke"]
    assert abs(e1 - e0) <= 5e-3 * abs(e0)
    assert ctx.neigh_stats()["nbuilds"] == 6  # setup + every 20 steps
    ctx.close()


@pytest.mark.parametrize("list_kind", ["half", "full"])
@pytest.mark.parametrize("mix", ["geometric", "arithmetic"])
def test_lj_mixed_cutoffs_gpu(pkg, mix, list_kind):
    """two types with different per-type cutoffs, the 1-2 cutoff mixed like PairLJCut::init_one
    (tests/golden/ljmix.npz, from the unmodified reference)"""
    from test_oracle_golden import ljmix_case
    B = pkg.b200md
    g, T, s, cn, cmax = ljmix_case(pkg, mix)
    ctx = B.Context(0)
    ctx.set_units("lj"); ctx.set_box(s["boxlo"], s["boxhi"])
    ctx.set_atoms(s["x"], s["v"], s["type"], s["tag"], s["mass"])
    ctx.pair_lj_cut_config(*[T[k] for k in ("cutsq", "lj1", "lj2", "lj3", "lj4", "offset")])
    ctx.neigh_config(cn, cmax, 0.3, every=20, delay=0, check=0,
                     list_kind=B.LIST_HALF_NEWTON if list_kind == "half" else B.LIST_FULL)
    ctx.setup()
    a = ctx.get_atoms()
    if list_kind == "half":
        nn, flat = ctx.neigh_export(by_tag=True)
        want = np.sort(g[mix + "_half_lo"].astype(np.int64) * (1 << 32) + g[mix + "_half_hi"].astype(np.int64))
        assert np.array_equal(Cm.pairs_from_lists(a["tag"], nn, flat), want)
    f = a["f"][np.argsort(a["tag"])]
    assert np.abs(f - g[mix + "_f"]).max() <= F_TOL * np.abs(g[mix + "_f"]).max()
    t = ctx.thermo()
    for k, key in enumerate(KEYS):
        assert abs(t[key] - g[mix + "_thermo0"][k]) <= T_TOL * abs(g[mix + "_thermo0"][k]), key
    t = ctx.run(20, 0.005, 0)[-1]
    for k, key in enumerate(KEYS):
        assert abs(t[key] - g[mix + "_thermo20"][k]) <= T_TOL * abs(g[mix + "_thermo20"][k]), key
    ctx.close()


@pytest.mark.parametrize("cluster", ["on", "off"])
@pytest.mark.parametrize("half", [True, False])
@pytest.mark.parametrize("name,kind", CASES)
def test_golden_both_list_layouts(pkg, oracle_mod, name, kind, half, cluster):
    """the cluster-pair list (4 atoms share one j-list, 4-bit pair masks; default for lj/cut) and the per-atom list
    (default for eam and molecular systems) must give the same exact per-atom lists, forces and thermo for every
    pair style: option `cluster` forces either layout"""
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    ctx, s = make_ctx(pkg, oracle_mod, g, kind, B.LIST_HALF_NEWTON if half else B.LIST_FULL, 4)
    ctx.set_option("cluster", cluster)
    assert ctx.get_option("cluster") == cluster
    ctx.setup()
    a = ctx.get_atoms()
    nn, flat = ctx.neigh_export(by_tag=True)
    if half:
        want = np.sort(g["half_lo"].astype(np.int64) * (1 << 32) + g["half_hi"].astype(np.int64))
        assert np.array_equal(Cm.pairs_from_lists(a["tag"], nn, flat), want)
    else:
        want = np.sort(g["full_i"].astype(np.int64) * (1 << 32) + g["full_j"].astype(np.int64))
        assert np.array_equal(Cm.full_sets(a["tag"], nn, flat), want)
    order = np.argsort(a["tag"])
    assert np.abs(a["f"][order] - g["f"]).max() <= F_TOL * np.abs(g["f"]).max()
    t = ctx.thermo()
    for k, key in enumerate(KEYS):
        assert abs(t[key] - g["thermo0"][k]) <= T_TOL * abs(g["thermo0"][k]), key
    for row in g["thermo_series"][:3]:
        ctx.setup()
        t = ctx.run(10, s["dt"], 0)[-1]
        for k, key in enumerate(KEYS):
            assert abs(t[key] - row[k]) <= T_TOL * abs(row[k]), (key, t[key], row[k])
    ctx.close()


@pytest.mark.parametrize("kind", ["lj", "eam"])
def test_graph_mode_is_the_same_run(pkg, kind):
    """option graph: the steps between two list builds replayed as one CUDA graph (a WHILE node whose condition is
    Neighbor::decide evaluated on the device) must be the SAME run as the host-driven loop: with the full list
    (newton off, no atomics) positions and velocities are bit-identical, and the rebuilds fall on the same steps
    (`check no` every 20 for in.lj; `delay 5 every 1 check yes` for in.eam, where the distance test decides)."""
    B = pkg.b200md
    res = {}
    for mode in ("off", "on"):
        ctx = B.Context(0)
        ctx.set_option("graph", mode)
        cfg = pkg.inputs.in_lj(ctx, list_kind=B.LIST_FULL) if kind == "lj" else pkg.inputs.in_eam(ctx, list_kind=B.LIST_FULL)
        ctx.setup()
        th = ctx.run(57, cfg["dt"], 25)          # output steps 25, 50 and 57 cut the graph spans
        th += ctx.run(43, cfg["dt"], 0)
        a = ctx.get_atoms(("x", "v", "tag"))
        o = np.argsort(a["tag"])
        res[mode] = (a["x"][o], a["v"][o], ctx.neigh_stats(), [(t["step"], t["temp"], t["pe"], t["press"]) for t in th])
        ctx.close()
    assert np.array_equal(res["on"][0], res["off"][0]) and np.array_equal(res["on"][1], res["off"][1])
    assert res["on"][2]["nbuilds"] == res["off"][2]["nbuilds"] and res["on"][2]["ndanger"] == res["off"][2]["ndanger"]
    assert res["on"][3] == res["off"][3]
    assert res["on"][2]["nbuilds"] == (6 if kind == "lj" else res["off"][2]["nbuilds"])


def test_streamed_offload_boundary_is_the_same_run(pkg):
    """The e2e boundary of bench.py: a host that takes x and f down and sends x back up every step.  Whole-array
    blocking copies (put_atoms / get_atoms), the asynchronous download (get_atoms_async / wait_download) and the
    streamed form (wait_download_part + put_atoms_range piece by piece) must be the SAME run bit for bit on the full
    list (no atomics), across a rebuild (the local order changes), and the pieces must tile the owned atoms."""
    B = pkg.b200md
    res = {}
    for mode in ("sync", "async", "stream"):
        ctx = B.Context(0)
        cfg = pkg.inputs.in_lj(ctx, 2, 2, 2, list_kind=B.LIST_FULL)      # 256 000 atoms: 3 pieces of x
        ctx.setup()
        n = ctx.nlocal()[0]
        hx, hf = B.pinned_empty((n, 3)), B.pinned_empty((n, 3))
        pieces = []
        for step in range(25):                                            # rebuild at step 20
            ctx.run(1, cfg["dt"], 0)
            if mode == "sync":
                ctx.get_atoms_into(hx, None, hf, None)
                ctx.put_atoms(hx, None)
            elif mode == "async":
                ctx.get_atoms_async(hx, hf)
                ctx.wait_download(False)
                ctx.put_atoms(hx, None)
                ctx.wait_download(True)
            else:
                ctx.get_atoms_async(hx, hf)
                p, covered = 0, 0
                while True:
                    first, count = ctx.wait_download_part(p)
                    if count == 0:
                        break
                    assert first == covered
                    ctx.put_atoms_range(hx, first, count)
                    covered += count
                    p += 1
                assert covered == n
                pieces.append(p)
                ctx.wait_download(True)
        a = ctx.get_atoms(("x", "v", "f", "tag"))
        o = np.argsort(a["tag"])
        res[mode] = (a["x"][o], a["v"][o], a["f"][o], hx.copy(), hf.copy(), ctx.neigh_stats()["nbuilds"])
        if mode == "stream":
            assert min(pieces) >= 2, pieces
            with pytest.raises(B.B200MDError, match="range outside"):
                ctx.put_atoms_range(hx, n - 10, 11)
        ctx.close()
    for mode in ("async", "stream"):
        for k in range(5):
            assert np.array_equal(res[mode][k], res["sync"][k]), (mode, k)
        assert res[mode][5] == res["sync"][5] == 2


@pytest.mark.parametrize("order", ["x", "index"])
@pytest.mark.parametrize("scale", ["0.5", "0.8", "1.0", "1.7", "3.0"])
@pytest.mark.parametrize("half", [True, False])
@pytest.mark.parametrize("name,kind", CASES)
def test_lists_do_not_depend_on_the_bins(pkg, oracle_mod, name, kind, half, scale, order):
    """options bin_scale / cell_order: the device picks its own bin size and in-bin order.  Membership is
    decided by distance and the half rule, so for every bin size — smaller and larger than the reference's
    cutneighmax/2, i.e. more and fewer stencil rows than NStencil builds — and either order the per-atom lists must
    stay the reference's exact sets, and forces / thermo over 30 steps (two or more rebuilds) the golden ones."""
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    ctx, s = make_ctx(pkg, oracle_mod, g, kind, B.LIST_HALF_NEWTON if half else B.LIST_FULL, 2)
    ctx.set_option("bin_scale", scale)
    ctx.set_option("cell_order", order)
    assert float(ctx.get_option("bin_scale")) == float(scale) and ctx.get_option("cell_order") == order
    ctx.setup()
    a = ctx.get_atoms()
    nn, flat = ctx.neigh_export(by_tag=True)
    if half:
        want = np.sort(g["half_lo"].astype(np.int64) * (1 << 32) + g["half_hi"].astype(np.int64))
        assert np.array_equal(Cm.pairs_from_lists(a["tag"], nn, flat), want)
    else:
        want = np.sort(g["full_i"].astype(np.int64) * (1 << 32) + g["full_j"].astype(np.int64))
        assert np.array_equal(Cm.full_sets(a["tag"], nn, flat), want)
    o = np.argsort(a["tag"])
    assert np.abs(a["f"][o] - g["f"]).max() <= F_TOL * np.abs(g["f"]).max()
    for row in g["thermo_series"][:3]:
        ctx.setup()
        t = ctx.run(10, s["dt"], 0)[-1]
        for k, key in enumerate(KEYS):
            assert abs(t[key] - row[k]) <= T_TOL * abs(row[k]), (key, t[key], row[k])
    ctx.close()


@pytest.mark.parametrize("name,kind", [c for c in CASES if c[1] != "lj"])
def test_eam_masks_and_shared_tables_change_nothing(pkg, oracle_mod, name, kind):
    """options eam_mask / eam_smem: the force pass walking only the pairs the density pass found inside the cutoff, and
    the one-element tables read from (swizzled) shared memory, are the same arithmetic on the same pairs — forces equal
    to rounding (the RED order differs), thermo over 30 steps equal to the golden run in every combination."""
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    ref = None
    for mask in ("on", "off"):
        for smem in ("auto", "on", "off"):
            ctx, s = make_ctx(pkg, oracle_mod, g, kind, B.LIST_HALF_NEWTON, 1)
            ctx.set_option("eam_mask", mask)
            ctx.set_option("eam_smem", smem)
            assert ctx.get_option("eam_mask") == mask and ctx.get_option("eam_smem") == smem
            ctx.setup()
            a = ctx.get_atoms(("f", "tag"))
            f = a["f"][np.argsort(a["tag"])]
            assert np.abs(f - g["f"]).max() <= F_TOL * np.abs(g["f"]).max()
            if ref is None:
                ref = f
            assert np.abs(f - ref).max() <= 1e-13 * np.abs(ref).max()
            for row in g["thermo_series"][:3]:
                ctx.setup()
                t = ctx.run(10, s["dt"], 0)[-1]
                for k, key in enumerate(KEYS):
                    assert abs(t[key] - row[k]) <= T_TOL * abs(row[k]), (mask, smem, key, t[key], row[k])
            ctx.close()


@pytest.mark.parametrize("mode", ["direct", "staged"])
@pytest.mark.parametrize("half", [True, False])
@pytest.mark.parametrize("name,kind", CASES)
def test_one_rank_border_modes(pkg, oracle_mod, name, kind, half, mode):
    """option self_border: the one-pass 26-direction ghost images (default) and the reference's three stages
    (CommBrick::borders, comm_brick.cpp:771-979) must produce the same ghost SET — same number of ghosts, same multiset
    of (tag, position) — and with it the reference's exact lists, forces and thermo."""
    B = pkg.b200md
    g = np.load(os.path.join(Cm.GOLDEN, name + ".npz"))
    res = {}
    for m in (mode, "staged"):
        ctx, s = make_ctx(pkg, oracle_mod, g, kind, B.LIST_HALF_NEWTON if half else B.LIST_FULL, 2)
        ctx.set_option("self_border", m)
        assert ctx.get_option("self_border") == m
        ctx.setup()
        gh = ctx.get_ghosts()
        key = np.lexsort((gh["x"][:, 2], gh["x"][:, 1], gh["x"][:, 0], gh["tag"]))
        a = ctx.get_atoms()
        nn, flat = ctx.neigh_export(by_tag=True)
        sets = Cm.pairs_from_lists(a["tag"], nn, flat) if half else Cm.full_sets(a["tag"], nn, flat)
        th = []
        for row in g["thermo_series"][:2]:
            ctx.setup()
            th.append(ctx.run(10, s["dt"], 0)[-1])
        res[m] = (gh["tag"][key], gh["x"][key], sets, a["f"][np.argsort(a["tag"])], th)
        ctx.close()
    want = (np.sort(g["half_lo"].astype(np.int64) * (1 << 32) + g["half_hi"].astype(np.int64)) if half
            else np.sort(g["full_i"].astype(np.int64) * (1 << 32) + g["full_j"].astype(np.int64)))
    assert np.array_equal(res[mode][2], want)
    assert np.array_equal(res[mode][0], res["staged"][0]) and np.array_equal(res[mode][1], res["staged"][1])
    assert np.abs(res[mode][3] - g["f"]).max() <= F_TOL * np.abs(g["f"]).max()
    for t, row in zip(res[mode][4], g["thermo_series"][:2]):
        for k, key_ in enumerate(KEYS):
            assert abs(t[key_] - row[k]) <= T_TOL * abs(row[k]), (key_, t[key_], row[k])
