"""Drop-in test: the reference program itself (its sources, compiled with the suffix styles of
lammps-kokkos-port_b200/lammps_styles registered -> `lmp_b200`) runs UNMODIFIED bench input
scripts with `-sf b200`, and its thermo output is compared with
  * the numbers printed in the reference's own logs (bench/log.*.lj.* / log.*.eam.*), and
  * the unmodified reference executable (oracle/_ref/lmp) on the same script, 12 digits.
The scripts below are the text of bench/in.lj and bench/in.eam (variables expanded)."""
import os
import re
import subprocess
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LMP_B200 = os.path.join(ROOT, "lammps-kokkos-port_b200", "lammps_styles", "_build", "lmp_b200")
LMP_REF = os.path.join(ROOT, "oracle", "_ref", "lmp")
POT = os.path.join(ROOT, "oracle", "_ref", "potentials")

IN_LJ = """
units           lj
atom_style      atomic
lattice         fcc 0.8442
region          box block 0 {n} 0 {n} 0 {n}
create_box      1 box
create_atoms    1 box
mass            1 1.0
velocity        all create 1.44 87287 loop geom
pair_style      lj/cut 2.5
pair_coeff      1 1 1.0 1.0 2.5
neighbor        0.3 bin
neigh_modify    delay 0 every 20 check no
fix             1 all nve
thermo_modify   format float %.15g
thermo          {thermo}
run             100
"""

IN_EAM = """
units           metal
atom_style      atomic
lattice         fcc 3.615
region          box block 0 {n} 0 {n} 0 {n}
create_box      1 box
create_atoms    1 box
pair_style      {style}
pair_coeff      {coeff}
velocity        all create 1600.0 376847 loop geom
neighbor        1.0 bin
neigh_modify    every 1 delay 5 check yes
fix             1 all nve
timestep        0.005
thermo_modify   format float %.15g
thermo          50
run             100
"""


def run_lmp(exe, script, extra=()):
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "in.script")
        with open(path, "w") as fh:
            fh.write(script)
        r = subprocess.run([exe, "-in", path, "-log", "none", *extra], cwd=d, capture_output=True, text=True,
                           timeout=900)
    return r.returncode, r.stdout + r.stderr


def thermo_rows(out):
    """rows between the 'Step Temp ...' header and 'Loop time'"""
    rows, on = [], False
    for line in out.splitlines():
        if re.match(r"\s*Step\s+Temp", line):
            on = True
            continue
        if line.startswith("Loop time"):
            on = False
        if on:
            try:
                rows.append([float(t) for t in line.split()])
            except ValueError:
                pass
    return np.array(rows)


def have_gpu_binary():
    if not os.path.exists(LMP_B200):
        pytest.fail("lmp_b200 missing: run __graft_entry__.build() where /root/reference exists")


@pytest.mark.gpu
def test_in_lj_matches_reference_log():
    """bench/in.lj (32000 atoms): log.*.lj.* print T/E_pair/Press at steps 0 and 100."""
    have_gpu_binary()
    rc, out = run_lmp(LMP_B200, IN_LJ.format(n=20, thermo=100), ("-sf", "b200"))
    assert rc == 0, out[-2000:]
    assert "Setting up Verlet/b200 run" in out
    rows = thermo_rows(out)
    assert rows.shape == (2, 6)
    # Step Temp E_pair E_mol TotEng Press (reference log, 8 significant digits printed)
    np.testing.assert_allclose(rows[0, 1:], [1.44, -6.7733681, 0.0, -4.6134356, -5.0197073], rtol=1e-7, atol=1e-12)
    np.testing.assert_allclose(rows[1, 1:], [0.7574531, -5.7585055, 0.0, -4.6223613, 0.20726105], rtol=1e-7, atol=1e-12)
    m = re.search(r"Neighbor list builds = (\d+)", out)
    assert m and int(m.group(1)) == 5
    assert "Dangerous builds not checked" in out


@pytest.mark.gpu
@pytest.mark.parametrize("n,thermo", [(10, 10), (20, 25)])
def test_in_lj_matches_reference_executable(n, thermo):
    have_gpu_binary()
    if not os.path.exists(LMP_REF):
        pytest.skip("oracle/_ref/lmp not built")
    script = IN_LJ.format(n=n, thermo=thermo)
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc == 0, out[-2000:]
    rc2, ref = run_lmp(LMP_REF, script)
    assert rc2 == 0, ref[-2000:]
    a, b = thermo_rows(out), thermo_rows(ref)
    assert a.shape == b.shape and a.shape[0] == 100 // thermo + 1
    np.testing.assert_allclose(a, b, rtol=1e-8, atol=1e-10)


@pytest.mark.gpu
@pytest.mark.parametrize("style,coeff", [("eam", "1 1 {pot}/Cu_u3.eam"),
                                         ("eam/alloy", "* * {pot}/Cu_mishin1.eam.alloy Cu")])
def test_in_eam_matches_reference_executable(style, coeff):
    have_gpu_binary()
    if not os.path.exists(LMP_REF):
        pytest.skip("oracle/_ref/lmp not built")
    script = IN_EAM.format(n=10, style=style, coeff=coeff.format(pot=POT))
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc == 0, out[-2000:]
    rc2, ref = run_lmp(LMP_REF, script)
    assert rc2 == 0, ref[-2000:]
    a, b = thermo_rows(out), thermo_rows(ref)
    assert a.shape == b.shape == (3, 6)
    np.testing.assert_allclose(a, b, rtol=1e-8, atol=1e-9)
    ma, mb = (re.search(r"Neighbor list builds = (\d+)", t) for t in (out, ref))
    assert ma and mb and ma.group(1) == mb.group(1)
    da, db = (re.search(r"Dangerous builds = (\d+)", t) for t in (out, ref))
    assert da and db and da.group(1) == db.group(1)


@pytest.mark.gpu
def test_in_eam_32000_matches_reference_log():
    """bench/in.eam (32000 atoms, Cu_u3): log.*.eam.* thermo at steps 0, 50, 100."""
    have_gpu_binary()
    script = IN_EAM.format(n=20, style="eam", coeff="1 1 %s/Cu_u3.eam" % POT)
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc == 0, out[-2000:]
    rows = thermo_rows(out)
    assert rows.shape == (3, 6)
    ref = np.array([[1600.0, -113280.0, 0.0, -106662.09, 18703.573],
                    [781.69049, -109873.35, 0.0, -106640.13, 52273.088],
                    [801.832, -109957.3, 0.0, -106640.77, 51322.821]])
    np.testing.assert_allclose(rows[:, 1:], ref, rtol=6e-8, atol=1e-12)  # 8 printed digits


IN_CHAIN = """
units           lj
atom_style      bond
special_bonds   fene
read_data       {data}
neighbor        0.4 bin
neigh_modify    every 1 delay 1
bond_style      fene
bond_coeff      1 30.0 1.5 1.0 1.0
pair_style      lj/cut 1.12
pair_modify     shift yes
pair_coeff      1 1 1.0 1.0 1.12
fix             1 all nve
thermo_modify   format float %.15g
thermo          20
timestep        0.012
run             60
run             20
"""


@pytest.mark.gpu
def test_in_chain_matches_reference_executable():
    """bench/in.chain without `fix langevin` (its RNG stream depends on the host's atom order):
    atom_style bond, special_bonds fene, bond_style fene/b200 + lj/cut/b200 through lmp_b200.
    Two `run`s: the second re-uploads atoms and topology from the host arrays the first left."""
    have_gpu_binary()
    data = os.path.join(ROOT, "oracle", "_ref", "data", "data.chain")
    if not (os.path.exists(LMP_REF) and os.path.exists(data)):
        pytest.skip("oracle/_ref not built")
    script = IN_CHAIN.format(data=data)
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc == 0, out[-2000:]
    rc2, ref = run_lmp(LMP_REF, script)
    assert rc2 == 0, ref[-2000:]
    a, b = thermo_rows(out), thermo_rows(ref)
    assert a.shape == b.shape == (6, 6)
    np.testing.assert_allclose(a, b, rtol=1e-8, atol=1e-10)
    # step 0 as printed in the reference's log (bench/log.6Oct16.chain.fixed.icc.1)
    np.testing.assert_allclose(a[0, 1:], [0.97029772, 0.44484087, 20.494523, 22.394765, 4.6721833], rtol=6e-8)
    ma, mb = (re.findall(r"Neighbor list builds = (\d+)", t) for t in (out, ref))
    assert ma == mb and len(ma) == 2


@pytest.mark.gpu
def test_in_chain_with_langevin_matches_reference_executable():
    """bench/in.chain AS SHIPPED (fix nve + fix langevin 1.0 1.0 10.0 904297) with -sf b200:
    fix langevin/b200 replays the reference's RanMars stream on the device in the reference's
    local atom order (Atom::sort at set-up), so the thermo equals the reference executable's and
    the published log (bench/log.6Oct16.chain.fixed.icc.1, step 100).  The second `run` checks
    that the generator carries on across runs like the reference's (one RanMars per fix)."""
    have_gpu_binary()
    data = os.path.join(ROOT, "oracle", "_ref", "data", "data.chain")
    if not (os.path.exists(LMP_REF) and os.path.exists(data)):
        pytest.skip("oracle/_ref not built")
    script = IN_CHAIN.format(data=data).replace(
        "fix             1 all nve", "fix             1 all nve\nfix             2 all langevin 1.0 1.0 10.0 904297").replace(
        "run             60", "run             100")
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc == 0, out[-2000:]
    rc2, ref = run_lmp(LMP_REF, script)
    assert rc2 == 0, ref[-2000:]
    a, b = thermo_rows(out), thermo_rows(ref)
    assert a.shape == b.shape == (8, 6)
    np.testing.assert_allclose(a, b, rtol=1e-8, atol=1e-10)
    np.testing.assert_allclose(a[5, 1:], [0.9729966, 0.4361122, 20.507698, 22.40326, 4.6548819], rtol=6e-8)
    ma, mb = (re.findall(r"Neighbor list builds = (\d+)", t) for t in (out, ref))
    assert ma == mb and len(ma) == 2


@pytest.mark.gpu
def test_langevin_options_the_device_does_not_cover_fail_loudly():
    have_gpu_binary()
    script = IN_LJ.format(n=5, thermo=100).replace(
        "fix             1 all nve", "fix             1 all nve\nfix 2 all langevin 1.0 1.0 1.0 4711 zero yes")
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc != 0 and "langevin/b200" in out


PERATOM = """
compute         pea all pe/atom
compute         sta all stress/atom NULL
compute         stv all stress/atom NULL virial
variable        wpe atom c_pea*id
variable        wxy atom c_sta[4]*id
compute         r1 all reduce sum c_pea c_sta[1] c_sta[2] c_sta[3] c_stv[4] c_stv[5] c_stv[6] v_wpe v_wxy
compute         r2 all reduce max c_pea c_sta[1] c_stv[6]
compute         r3 all reduce min c_pea c_sta[1] c_stv[6]
thermo_style    custom step temp pe press c_r1[*] c_r2[*] c_r3[*]
"""


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["lj", "eam", "chain", "chain_newton_off"])
def test_peratom_computes_match_reference_executable(case):
    """compute pe/atom and stress/atom on top of the b200 styles (src/compute_pe_atom.cpp:92-170,
    src/compute_stress_atom.cpp:140-340 read Pair::eatom/vatom and Bond::eatom/vatom): sums, extrema
    and tag-weighted sums (which catch a permutation of the rows) against the reference executable."""
    have_gpu_binary()
    data = os.path.join(ROOT, "oracle", "_ref", "data", "data.chain")
    if not (os.path.exists(LMP_REF) and os.path.exists(data)):
        pytest.skip("oracle/_ref not built")
    if case == "lj":
        script = IN_LJ.format(n=6, thermo=20).replace("thermo_modify", PERATOM + "thermo_modify").replace("run             100", "run 40")
    elif case == "eam":
        script = IN_EAM.format(n=5, style="eam", coeff="1 1 %s/Cu_u3.eam" % POT).replace("thermo_modify", PERATOM + "thermo_modify").replace(
            "thermo          50", "thermo 20").replace("run             100", "run 40")
    else:
        script = IN_CHAIN.format(data=data).replace("thermo_modify", PERATOM + "thermo_modify").replace(
            "run             60", "run 40").replace("run             20\n", "")
        if case == "chain_newton_off":
            script = script.replace("special_bonds   fene", "special_bonds   fene\nnewton off on")
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc == 0, out[-2000:]
    rc2, ref = run_lmp(LMP_REF, script)
    assert rc2 == 0, ref[-2000:]
    a, b = thermo_rows(out), thermo_rows(ref)
    assert a.shape == b.shape == (3, 19)
    np.testing.assert_allclose(a, b, rtol=2e-8, atol=1e-8 * np.abs(b).max())


@pytest.mark.gpu
def test_unsupported_setups_fail_loudly():
    """no silent host fallback: a fix the device loop does not cover is an error"""
    have_gpu_binary()
    script = IN_LJ.format(n=5, thermo=100).replace("fix             1 all nve",
                                                   "fix             1 all nve\nfix 2 all momentum 10 linear 1 1 1")
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc != 0 and "verlet/b200" in out


def test_lmp_b200_without_gpu_fails_loudly():
    """CPU box: the binary exists (when it was built here) and refuses to run without a device."""
    if not os.path.exists(LMP_B200):
        pytest.skip("lmp_b200 not built in this checkout")
    rc, out = run_lmp(LMP_B200, IN_LJ.format(n=5, thermo=100), ("-sf", "b200"))
    if rc == 0:
        pytest.skip("a CUDA device is present")
    assert "no CUDA device" in out


IN_LISTS = """
units           lj
{newton}
atom_style      atomic
lattice         fcc 0.8442
region          box block 0 8 0 8 0 8
create_box      1 box
create_atoms    1 box
mass            1 1.0
velocity        all create 1.44 87287 loop geom
pair_style      lj/cut 2.5
pair_coeff      1 1 1.0 1.0 2.5
neighbor        0.3 bin
neigh_modify    delay 0 every 20 check no
fix             1 all nve
compute         c all coord/atom cutoff 1.3
compute         csum all reduce sum c_c
compute         cmax all reduce max c_c
compute         cmin all reduce min c_c
compute         r all rdf 50
thermo_style    custom step temp pe press c_csum c_cmax c_cmin c_r[12][2] c_r[23][2] c_r[23][3] c_r[50][3]
thermo_modify   format float %.15g
thermo          {thermo}
run             60
"""


@pytest.mark.gpu
@pytest.mark.parametrize("newton,thermo", [("", 10), ("", 7), ("newton off", 10)])
def test_host_neighbor_lists_are_published(newton, thermo):
    """NPairFullBinB200 / NPairHalfBinNewtonB200 (lammps_styles/npair_b200.cpp): host-side consumers of a neighbor
    list under `-sf b200` — compute coord/atom (an occasional FULL list) and compute rdf (an occasional half list, with
    newton on a copy of the pair style's own) — must print what they print under the unmodified reference executable.
    thermo 7 puts output steps between the rebuilds (every 20), where the device ghosts and the host ghosts were
    selected at different times."""
    have_gpu_binary()
    if not os.path.exists(LMP_REF):
        pytest.skip("oracle/_ref/lmp not built")
    script = IN_LISTS.format(newton=newton, thermo=thermo)
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc == 0, out[-3000:]
    rc2, ref = run_lmp(LMP_REF, script)
    assert rc2 == 0, ref[-2000:]
    a, b = thermo_rows(out), thermo_rows(ref)
    assert a.shape == b.shape and a.shape[1] == 11 and a.shape[0] >= 7
    np.testing.assert_allclose(a[:, :4], b[:, :4], rtol=1e-8, atol=1e-10)
    assert np.array_equal(a[:, 4:7], b[:, 4:7])                      # coordination numbers are integers
    np.testing.assert_allclose(a[:, 7:], b[:, 7:], rtol=1e-10, atol=1e-12)
    assert b[0, 4] > 0 and b[-1, 10] > 0


@pytest.mark.gpu
def test_fix_ave_time_is_served():
    """an end_of_step fix that only reads (fix ave/time over compute coord/atom's sum, the pair energy and the
    pressure): verlet/b200 stops the device loop on the steps it samples, refreshes the host arrays, tallies and
    lists, and runs Modify::end_of_step — same averages as the reference executable"""
    have_gpu_binary()
    if not os.path.exists(LMP_REF):
        pytest.skip("oracle/_ref/lmp not built")
    script = IN_LISTS.format(newton="", thermo=12).replace(
        "thermo_style    custom step temp pe press c_csum c_cmax c_cmin c_r[12][2] c_r[23][2] c_r[23][3] c_r[50][3]",
        "fix             av all ave/time 3 4 12 c_csum c_thermo_pe c_thermo_press\n"
        "thermo_style    custom step temp pe press f_av[1] f_av[2] f_av[3]")
    rc, out = run_lmp(LMP_B200, script, ("-sf", "b200"))
    assert rc == 0, out[-3000:]
    rc2, ref = run_lmp(LMP_REF, script)
    assert rc2 == 0, ref[-2000:]
    a, b = thermo_rows(out), thermo_rows(ref)
    assert a.shape == b.shape and a.shape[1] == 7
    np.testing.assert_allclose(a, b, rtol=1e-8, atol=1e-9)
    assert np.abs(b[-1, 4:]).min() > 0
