#!/usr/bin/env python3
"""Turn the reference's own pair-style fixtures (unittest/force-styles/tests/atomic-pair-eam*.yaml
+ in.metal + data.metal) into small .npz files the GPU box can read (it has no /root/reference).

What the harness does for these numbers (unittest/force-styles/test_pair_style.cpp:108-170):
  in.metal settings (atom_style atomic, neigh_modify delay 2 every 2 check no, timestep 1e-4,
  newton on), read_data data.metal (32 atoms, 2 types, 7 A periodic box), pair_style/pair_coeff
  from the YAML, `run 0 post no`  -> init_forces / init_vdwl / init_stress (= Pair::virial);
  `fix 1 all nve`, `run 4 post no` -> run_forces / run_vdwl / run_stress.
Run here (needs /root/reference):  python tests/golden/make_yaml_fixtures.py
"""
import os
import re

import numpy as np
import yaml

REF = os.environ.get("B200MD_REFERENCE", "/root/reference")
TDIR = os.path.join(REF, "unittest", "force-styles", "tests")
HERE = os.path.dirname(os.path.abspath(__file__))

CASES = ["atomic-pair-eam", "atomic-pair-eam_real", "atomic-pair-eam_alloy", "atomic-pair-eam_alloy_real",
         "atomic-pair-eam_fs", "atomic-pair-eam_fs_real"]


def read_data_metal(path):
    lines = open(path).read().splitlines()
    box = {}
    for ln in lines:
        m = re.match(r"\s*(\S+)\s+(\S+)\s+([xyz])lo \3hi", ln)
        if m:
            box[m.group(3)] = (float(m.group(1)), float(m.group(2)))
    def section(name, ncol):
        i = next(k for k, ln in enumerate(lines) if ln.split("#")[0].strip() == name) + 2
        rows = []
        while i < len(lines) and lines[i].strip():
            rows.append(lines[i].split()[:ncol])
            i += 1
        return rows
    masses = section("Masses", 2)
    atoms = section("Atoms", 8)
    vels = section("Velocities", 4)
    n = len(atoms)
    x = np.zeros((n, 3)); v = np.zeros((n, 3)); typ = np.zeros(n, np.int32); img = np.zeros((n, 3), np.int32)
    for r in atoms:
        i = int(r[0]) - 1
        typ[i] = int(r[1]); x[i] = [float(t) for t in r[2:5]]; img[i] = [int(t) for t in r[5:8]]
    for r in vels:
        v[int(r[0]) - 1] = [float(t) for t in r[1:4]]
    mass = np.zeros(len(masses) + 1)
    for r in masses:
        mass[int(r[0])] = float(r[1])
    return dict(boxlo=np.array([box[k][0] for k in "xyz"]), boxhi=np.array([box[k][1] for k in "xyz"]),
                x=x, v=v, type=typ, image=img, mass=mass)


def table(text, ncol):
    rows = [ln.split() for ln in text.strip().splitlines()]
    a = np.array([[float(t) for t in r] for r in rows])
    assert a.shape[1] == ncol
    return a


def main():
    d = read_data_metal(os.path.join(TDIR, "data.metal"))
    for case in CASES:
        y = yaml.safe_load(open(os.path.join(TDIR, case + ".yaml")))
        units = "real" if "units index real" in (y.get("pre_commands") or "") else "metal"
        fi = table(y["init_forces"], 4); fr = table(y["run_forces"], 4)
        assert np.array_equal(fi[:, 0], np.arange(1, 33)) and np.array_equal(fr[:, 0], np.arange(1, 33))
        out = dict(d)
        out.update(units=units, pair_style=y["pair_style"], pair_coeff=np.array(y["pair_coeff"].strip().splitlines()),
                   epsilon=float(y["epsilon"]), init_vdwl=float(y["init_vdwl"]), run_vdwl=float(y["run_vdwl"]),
                   init_stress=np.array([float(t) for t in y["init_stress"].split()]),
                   run_stress=np.array([float(t) for t in y["run_stress"].split()]),
                   init_forces=fi[:, 1:], run_forces=fr[:, 1:])
        np.savez_compressed(os.path.join(HERE, "yaml_" + case.replace("atomic-pair-", "") + ".npz"), **out)
        print(case, units, y["pair_style"], list(out["pair_coeff"]))


if __name__ == "__main__":
    main()
