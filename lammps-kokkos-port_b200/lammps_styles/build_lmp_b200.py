#!/usr/bin/env python3
"""Build `lmp_b200`: the reference program (LAMMPS 2Aug2023 fork, sources compiled where they
lie under /root/reference/src, nothing copied) with the suffix styles of this directory
registered, linked against libb200md.so.  This is the drop-in proof: an unmodified input
script run with `-sf b200` executes the hot path on the device.

The reference's own build systems are not run; this script writes a ninja file.  The style
registry headers (style_pair.h, style_fix.h, style_integrate.h, ...) are generated the way the
reference's `Make.sh style` would: one #include line per selected style header -- here the
core + MANYBODY eam headers plus the *_b200.h headers of this directory.

Outputs under lammps_styles/_build/ (git-ignored; the binary ships to the GPU box):
  lmp_b200
Serial MPI stub (src/STUBS): one rank.  With a real MPI the same sources build unchanged.
"""
import glob
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
ROOT = os.path.dirname(PKG)
REF = os.environ.get("B200MD_REFERENCE", "/root/reference")
SRC = os.path.join(REF, "src")
OUT = os.path.join(HERE, "_build")
GEN = os.path.join(OUT, "gen")
OBJ = os.path.join(OUT, "obj")
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"

STYLES = [
    ("ANGLE_CLASS", "angle_", "angle"), ("ATOM_CLASS", "atom_vec_", "atom"),
    ("BODY_CLASS", "body_", "body"), ("BOND_CLASS", "bond_", "bond"),
    ("COMMAND_CLASS", "", "command"), ("COMPUTE_CLASS", "compute_", "compute"),
    ("DIHEDRAL_CLASS", "dihedral_", "dihedral"), ("DUMP_CLASS", "dump_", "dump"),
    ("FIX_CLASS", "fix_", "fix"), ("IMPROPER_CLASS", "improper_", "improper"),
    ("INTEGRATE_CLASS", "", "integrate"), ("KSPACE_CLASS", "", "kspace"),
    ("MINIMIZE_CLASS", "min_", "minimize"), ("NBIN_CLASS", "nbin_", "nbin"),
    ("NPAIR_CLASS", "npair_", "npair"), ("NSTENCIL_CLASS", "nstencil_", "nstencil"),
    ("NTOPO_CLASS", "ntopo_", "ntopo"), ("PAIR_CLASS", "pair_", "pair"),
    ("READER_CLASS", "reader_", "reader"), ("REGION_CLASS", "region_", "region"),
]


def write_if_changed(path, text):
    if os.path.exists(path) and open(path).read() == text:
        return
    with open(path, "w") as fh:
        fh.write(text)


def main():
    exe = os.path.join(OUT, "lmp_b200")
    if not os.path.isdir(SRC):
        print("build_lmp_b200: %s absent -- keeping prebuilt %s as is" % (REF, exe))
        return 0
    os.makedirs(GEN, exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    ref_srcs = sorted(glob.glob(os.path.join(SRC, "*.cpp")))
    ref_srcs += [os.path.join(SRC, "STUBS", "mpi.cpp")]
    ref_srcs += [os.path.join(SRC, "MANYBODY", f) for f in ("pair_eam.cpp", "pair_eam_alloy.cpp", "pair_eam_fs.cpp")]
    ref_srcs += [os.path.join(SRC, "MOLECULE", f) for f in ("atom_vec_bond.cpp", "bond_fene.cpp")]
    own_srcs = sorted(glob.glob(os.path.join(HERE, "*_b200.cpp")))
    headers = [s[:-4] + ".h" for s in ref_srcs + own_srcs if os.path.exists(s[:-4] + ".h")]
    for macro, prefix, name in STYLES:
        lines = []
        for h in sorted(headers, key=os.path.basename):
            base = os.path.basename(h)
            if prefix and not base.startswith(prefix):
                continue
            if re.search(r"^#ifdef\s+" + macro + r"\b", open(h).read(), re.M):
                lines.append('#include "%s"\n' % base)
        write_if_changed(os.path.join(GEN, "style_%s.h" % name), "".join(lines))
        write_if_changed(os.path.join(GEN, "packages_%s.h" % name), "")
    write_if_changed(os.path.join(GEN, "lmpinstalledpkgs.h"),
                     'const char * LAMMPS_NS::LAMMPS::installed_packages[] = '
                     '{"MANYBODY", "MOLECULE", "B200", NULL};\n')
    write_if_changed(os.path.join(GEN, "lmpgitversion.h"),
                     'bool LAMMPS_NS::LAMMPS::has_git_info() { return false; }\n'
                     'const char *LAMMPS_NS::LAMMPS::git_commit() { return "(unknown)"; }\n'
                     'const char *LAMMPS_NS::LAMMPS::git_branch() { return "(unknown)"; }\n'
                     'const char *LAMMPS_NS::LAMMPS::git_descriptor() { return "(unknown)"; }\n')
    inc = ["-I" + GEN, "-I" + SRC, "-I" + os.path.join(SRC, "STUBS"), "-I" + os.path.join(SRC, "MANYBODY"),
           "-I" + os.path.join(SRC, "MOLECULE"), "-I" + HERE, "-I" + os.path.join(ROOT, "include")]
    flags = ["-O2", "-DNDEBUG", "-ffp-contract=off", "-fPIC", "-std=c++11", "-DLAMMPS_SMALLBIG",
             "-DLAMMPS_EXCEPTIONS", "-w"]
    nj = ["cxx = %s" % CXX, "flags = %s %s" % (" ".join(flags), " ".join(inc)),
          "rule cc", "  command = $cxx $flags -MMD -MF $out.d -c $in -o $out",
          "  depfile = $out.d", "  deps = gcc", "  description = CC $in",
          "rule exe", "  description = LINK $out",
          "  command = $cxx -o $out $in -L%s -lb200md -Wl,-rpath,'$$ORIGIN/../..' -Wl,--disable-new-dtags" % PKG,
          ""]
    objs = []
    for s in ref_srcs + own_srcs:
        base = os.path.relpath(s, SRC if s.startswith(SRC) else HERE).replace("/", "_")[:-4] + ".o"
        o = os.path.join(OBJ, base)
        nj.append("build %s: cc %s" % (o, s))
        objs.append(o)
    nj.append("build %s: exe %s | %s" % (exe, " ".join(objs), os.path.join(PKG, "libb200md.so")))
    nj.append("default %s" % exe)
    write_if_changed(os.path.join(OUT, "build.ninja"), "\n".join(nj) + "\n")
    r = subprocess.run(["ninja", "-C", OUT, "-j", str(os.cpu_count() or 4)])
    if r.returncode == 0:
        print("build_lmp_b200: ok ->", exe)
    return r.returncode


if __name__ == "__main__":
    sys.exit(main())
