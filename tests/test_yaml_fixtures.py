"""The reference's own pair-style fixtures for this path
(unittest/force-styles/tests/atomic-pair-eam*.yaml: 32 atoms, 2 types, 7 A periodic box, forces /
energy / Pair::virial after `run 0` and after 4 NVE steps; epsilon 5e-12..7.5e-12), committed as
tests/golden/yaml_*.npz by tests/golden/make_yaml_fixtures.py.

The box is smaller than the ghost cutoff (2 periodic images per side), types differ, eam/fs has
asymmetric densities and the *_real cases exercise the readers' unit conversion: the multi-type
generality row of SURVEY.md section 8(f).  Three consumers:
  * CPU   : the C oracle with tables from eam_host.py
  * GPU   : the CUDA path through the C ABI (ctypes) with the same tables
  * GPU   : lmp_b200 -sf b200 -- the reference's readers feed the CUDA path
Tolerance: the harness compares with epsilon * (a few) relative to the LARGEST component; here
|got - want| <= 50 * epsilon * max|want| (forces) -- the same yardstick the reference applies to
its own accelerated variants (test_pair_style.cpp: EXPECT_FORCES with 10*epsilon for omp/gpu)."""
import os
import re
import subprocess
import tempfile

import numpy as np
import pytest

import common as Cm

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LMP_B200 = os.path.join(ROOT, "lammps-kokkos-port_b200", "lammps_styles", "_build", "lmp_b200")
POT = os.path.join(ROOT, "oracle", "_ref", "potentials")
CASES = ["eam", "eam_real", "eam_alloy", "eam_alloy_real", "eam_fs", "eam_fs_real"]
STEPS, DT = 4, 0.0001            # in.metal: timestep 0.0001; harness: run 4
EVERY, DELAY, CHECK = 2, 2, 0    # in.metal: neigh_modify delay 2 every 2 check no
SKIN = 2.0


def load(case):
    return np.load(os.path.join(Cm.GOLDEN, "yaml_%s.npz" % case))


def tables(pkg, g):
    E = pkg.eam_host
    style, units = str(g["pair_style"]), str(g["units"])
    coeff = [str(c).split() for c in g["pair_coeff"]]
    if style == "eam":
        files = [None, None]
        for c in coeff:
            files[int(c[0]) - 1] = os.path.join(POT, c[2])
        return E.tables_from_funcfl(files, units)
    if style == "eam/alloy":
        return E.tables_from_setfl(os.path.join(POT, coeff[0][2]), coeff[0][3:], units)
    return E.tables_from_fs(os.path.join(POT, coeff[0][2]), coeff[0][3:], units)


def check(g, f, vdwl, virial, which, scale=50.0):
    eps = float(g["epsilon"]) * scale
    want = g[which + "_forces"]
    assert np.abs(f - want).max() <= eps * np.abs(want).max(), (which, np.abs(f - want).max())
    assert abs(vdwl - float(g[which + "_vdwl"])) <= eps * abs(float(g[which + "_vdwl"]))
    ws = g[which + "_stress"]
    assert np.abs(np.asarray(virial) - ws).max() <= eps * np.abs(ws).max(), (which, virial, ws)


def by_tag(tag, arr):
    out = np.empty_like(arr)
    out[tag - 1] = arr
    return out


def cutneigh(pkg, T, ntypes):
    cutsq = np.zeros((ntypes + 1, ntypes + 1)); cutsq[1:, 1:] = T["cutforcesq"]
    return pkg.b200md.cutneighsq_table(ntypes, cutsq, SKIN)


@pytest.mark.parametrize("case", CASES)
def test_oracle_matches_yaml(pkg, oracle_mod, case):
    g = load(case)
    T = tables(pkg, g)
    n = len(g["type"])
    U = pkg.b200md.UNITS[str(g["units"])]
    o = oracle_mod.Oracle(g["boxlo"], g["boxhi"], g["x"], g["v"], g["type"], np.arange(1, n + 1, dtype=np.int32),
                          T["mass"], U)   # pair_coeff sets the masses from the potential file
    o.pair_eam(T)
    cn, cmax = cutneigh(pkg, T, 2)
    o.neigh_config(cn, cmax, SKIN, EVERY, DELAY, CHECK, True)
    o.setup()
    a = o.atoms(); th = o.thermo()
    check(g, by_tag(a["tag"], a["f"]), th["pe"], th["virial"], "init", scale=10.0)
    th = o.run(STEPS, DT, 0)[-1]
    a = o.atoms()
    check(g, by_tag(a["tag"], a["f"]), th["pe"], th["virial"], "run", scale=10.0)
    o.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
@pytest.mark.parametrize("lanes,half", [(4, False), (32, False), (8, True), (1, False), (2, True)])
def test_cuda_matches_yaml(pkg, case, lanes, half):
    B = pkg.b200md
    g = load(case)
    T = tables(pkg, g)
    n = len(g["type"])
    ctx = B.Context(0)
    ctx.set_lanes_per_atom(lanes)
    ctx.set_units(str(g["units"])); ctx.set_box(g["boxlo"], g["boxhi"])
    ctx.set_atoms(g["x"], g["v"], g["type"], np.arange(1, n + 1, dtype=np.int32), T["mass"])
    ctx.pair_eam_config(T)
    cn, cmax = cutneigh(pkg, T, 2)
    ctx.neigh_config(cn, cmax, SKIN, every=EVERY, delay=DELAY, check=CHECK, list_kind=B.LIST_HALF_NEWTON if half else B.LIST_FULL)
    ctx.setup()
    a = ctx.get_atoms(("tag", "f")); e, vir = ctx.pair_tallies()
    check(g, by_tag(a["tag"], a["f"]), e, vir, "init")
    ctx.run(STEPS, DT, 0)
    a = ctx.get_atoms(("tag", "f")); e, vir = ctx.pair_tallies()
    check(g, by_tag(a["tag"], a["f"]), e, vir, "run")
    ctx.close()


SCRIPT = """
units           {units}
atom_style      atomic
atom_modify     map array
neigh_modify    delay 2 every 2 check no
timestep        0.0001
newton          on on
pair_style      {style}
read_data       data.metal
{coeffs}
compute         vir all pressure NULL pair
variable        vol equal vol
thermo_style    custom step pe c_vir[1] c_vir[2] c_vir[3] c_vir[4] c_vir[5] c_vir[6] v_vol
thermo_modify   format float %.17g
run             0 post no
write_dump      all custom f0.txt id fx fy fz modify format float %.17g sort id
fix             1 all nve
thermo          2
run             4 post no
write_dump      all custom f4.txt id fx fy fz modify format float %.17g sort id
"""


def write_data(path, g):
    n = len(g["type"])
    with open(path, "w") as fh:
        fh.write("fixture\n\n%d atoms\n2 atom types\n\n" % n)
        for d, k in enumerate("xyz"):
            fh.write("%.17g %.17g %slo %shi\n" % (g["boxlo"][d], g["boxhi"][d], k, k))
        fh.write("\nMasses\n\n1 %.17g\n2 %.17g\n\nAtoms # atomic\n\n" % (g["mass"][1], g["mass"][2]))
        for i in range(n):
            fh.write("%d %d %.17g %.17g %.17g %d %d %d\n" % (i + 1, g["type"][i], *g["x"][i], *g["image"][i]))
        fh.write("\nVelocities\n\n")
        for i in range(n):
            fh.write("%d %.17g %.17g %.17g\n" % (i + 1, *g["v"][i]))


def read_dump(path):
    lines = open(path).read().splitlines()
    k = next(i for i, ln in enumerate(lines) if ln.startswith("ITEM: ATOMS"))
    return np.array([[float(t) for t in ln.split()[1:]] for ln in lines[k + 1:]])


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES)
def test_lmp_b200_matches_yaml(case):
    if not os.path.exists(LMP_B200):
        pytest.fail("lmp_b200 missing: run __graft_entry__.build() where /root/reference exists")
    g = load(case)
    coeffs = []
    for c in g["pair_coeff"]:
        coeffs.append("pair_coeff " + str(c))
    nktv2p = {"metal": 1.6021765e6, "real": 68568.415}[str(g["units"])]
    with tempfile.TemporaryDirectory() as d:
        write_data(os.path.join(d, "data.metal"), g)
        for c in g["pair_coeff"]:
            # the readers convert units only for files tagged "UNITS: metal" (potentials/*); the
            # bench/ copy of Cu_u3.eam has the same numbers without the tag
            name = str(c).split()[2]
            lines = open(os.path.join(POT, name)).read().split("\n")
            if "UNITS:" not in lines[0]:
                lines[0] = "UNITS: metal " + lines[0]
            with open(os.path.join(d, name), "w") as fh:
                fh.write("\n".join(lines))
        with open(os.path.join(d, "in.script"), "w") as fh:
            fh.write(SCRIPT.format(units=str(g["units"]), style=str(g["pair_style"]), coeffs="\n".join(coeffs)))
        r = subprocess.run([LMP_B200, "-sf", "b200", "-in", "in.script", "-log", "none"], cwd=d, capture_output=True,
                           text=True, timeout=600)
        out = r.stdout + r.stderr
        assert r.returncode == 0, out[-3000:]
        f0, f4 = read_dump(os.path.join(d, "f0.txt")), read_dump(os.path.join(d, "f4.txt"))
    rows = [[float(t) for t in ln.split()] for ln in out.splitlines() if re.match(r"\s+\d+\s+[-\d.eE+]+\s", ln)
            and len(ln.split()) == 9]
    rows = {int(r[0]): r for r in rows}
    for which, f, step in (("init", f0, 0), ("run", f4, 4)):
        r = rows[step]
        virial = np.array(r[2:8]) * r[8] / nktv2p      # P_ab = virial_ab / vol * nktv2p
        check(g, f, r[1], virial, which)
