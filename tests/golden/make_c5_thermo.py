"""Reference thermo of config C5 (bench/in.eam with `pair_style eam/alloy`,
`pair_coeff * * Cu_mishin1.eam.alloy Cu`, -var x/y/z 5 = 4 000 000 atoms), 100 steps, from the UNMODIFIED
reference executable compiled by oracle/build_ref.py (OPENMP package on all host cores; its thermo is
decomposition-independent to the printed digits).  Writes tests/golden/c5_thermo.json, which bench.py's
parity check and tests/test_gpu_parity.py compare the CUDA path with (8 digits, like the bench logs).

    python tests/golden/make_c5_thermo.py          # ~3 min on 8 cores, ~4 GB of host memory
"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as O  # noqa: E402


def main():
    x = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    body = O.IN_EAM.format(nx=20 * x, ny=20 * x, nz=20 * x, style="eam/alloy",
                           coeff="* * " + os.path.join(O.POT_DIR, "Cu_mishin1.eam.alloy") + " Cu")
    body = body.replace("atom_modify map array\n", "")
    body += "thermo 50\nthermo_style custom step temp pe ke press\nthermo_modify format float %.15g norm no\nrun 100\n"
    cores = os.cpu_count() or 1
    _, atoms, _, out = O.ref_exe_run(body, threads=cores, omp=True, timeout=3000)
    rows = {}
    for m in re.finditer(r"^\s*(\d+)\s+(\S+)\s+(\S+)\s+(\S+)\s+(\S+)\s*$", out, re.M):
        rows[int(m.group(1))] = dict(temp=float(m.group(2)), pe=float(m.group(3)), ke=float(m.group(4)),
                                     press=float(m.group(5)))
    assert set(rows) >= {0, 50, 100}, out[-3000:]
    nb = re.search(r"Neighbor list builds = (\d+)", out)
    dg = re.search(r"Dangerous builds = (\d+)", out)
    res = dict(config="in.eam, pair_style eam/alloy Cu_mishin1.eam.alloy, -var x %d -var y %d -var z %d" % (x, x, x),
               atoms=atoms, thermo={str(k): rows[k] for k in (0, 50, 100)},
               neighbor_list_builds=int(nb.group(1)), dangerous_builds=int(dg.group(1)),
               source="oracle/_ref/lmp -sf omp -pk omp %d (unmodified reference)" % cores)
    name = "c5_thermo.json" if x == 5 else "c5_thermo_x%d.json" % x
    with open(os.path.join(ROOT, "tests", "golden", name), "w") as fh:
        json.dump(res, fh, indent=1)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
