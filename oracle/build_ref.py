#!/usr/bin/env python3
"""Build the UNMODIFIED reference (LAMMPS 2Aug2023 fork at /root/reference) hot path into
oracle/_ref/ — TEST INFRASTRUCTURE ONLY (the checker and the timed CPU baseline).

Nothing of the reference is copied into the repo: sources are compiled where they lie
under /root/reference/src with /usr/bin/g++ by a ninja file this script writes.  The
reference's own build systems (cmake/, src/Makefile + Make.sh) are NOT run.  The only files
generated are the style-registry include lists (style_*.h = one `#include "x.h"` line per
style header that was selected, the same thing the reference's `Make.sh style` would list)
and two two-line version stubs; they live under oracle/_ref/gen/.

Selected sources:
  * every core file src/*.cpp (the serial MPI stub src/STUBS/mpi.cpp replaces MPI)
  * MANYBODY: pair_eam.cpp pair_eam_alloy.cpp pair_eam_fs.cpp   (the hot path's EAM styles)
  * MOLECULE: atom_vec_bond.cpp bond_fene.cpp                     (bench/in.chain)
  * OPENMP : the threaded variants of exactly the hot-path styles, used as the timed
             CPU baseline (`-sf omp -pk omp N`)

Outputs (all under oracle/_ref/, git-ignored, shipped to the GPU box by gpurun):
  liblammps_ref.so   reference as a shared library (C API of src/library.h, via ctypes)
  lmp                reference executable (src/main.cpp)
  potentials/, data/ Cu_u3.eam, Cu_mishin1.eam.alloy, ..., data.chain   (data the bench inputs name)

Flags: -O3 -DNDEBUG -ffp-contract=off (FMA-free like the survey's build),
-fopenmp, -DLAMMPS_SMALLBIG -DLAMMPS_EXCEPTIONS -DLMP_OPENMP.
"""
import glob
import os
import re
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("B200MD_REFERENCE", "/root/reference")
SRC = os.path.join(REF, "src")
OUT = os.path.join(HERE, "_ref")
GEN = os.path.join(OUT, "gen")
OBJ = os.path.join(OUT, "obj")
CXX = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"

MANYBODY = ["pair_eam.cpp", "pair_eam_alloy.cpp", "pair_eam_fs.cpp"]
MOLECULE = ["atom_vec_bond.cpp", "bond_fene.cpp"]      # bench/in.chain
OPENMP = [
    "bond_fene_omp.cpp",
    "fix_omp.cpp", "thr_omp.cpp", "thr_data.cpp", "domain_omp.cpp",
    "pair_lj_cut_omp.cpp", "pair_eam_omp.cpp", "pair_eam_alloy_omp.cpp",
    "fix_nve_omp.cpp",
    "npair_half_bin_atomonly_newton_omp.cpp", "npair_full_bin_atomonly_omp.cpp",
    "npair_half_bin_newton_omp.cpp", "npair_full_bin_omp.cpp",
]

# style-registry macro -> (header prefix the reference's Make.sh greps, style file name)
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


def selected_sources():
    srcs = sorted(glob.glob(os.path.join(SRC, "*.cpp")))
    srcs += [os.path.join(SRC, "STUBS", "mpi.cpp")]
    srcs += [os.path.join(SRC, "MANYBODY", f) for f in MANYBODY]
    srcs += [os.path.join(SRC, "MOLECULE", f) for f in MOLECULE]
    srcs += [os.path.join(SRC, "OPENMP", f) for f in OPENMP]
    return srcs


def write_if_changed(path, text):
    if os.path.exists(path) and open(path).read() == text:
        return
    with open(path, "w") as fh:
        fh.write(text)


def gen_headers(srcs):
    os.makedirs(GEN, exist_ok=True)
    headers = []
    for s in srcs:
        h = s[:-4] + ".h"
        if os.path.exists(h):
            headers.append(h)
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
                     '{"MANYBODY", "MOLECULE", "OPENMP", NULL};\n')
    write_if_changed(os.path.join(GEN, "lmpgitversion.h"),
                     'bool LAMMPS_NS::LAMMPS::has_git_info() { return false; }\n'
                     'const char *LAMMPS_NS::LAMMPS::git_commit() { return "(unknown)"; }\n'
                     'const char *LAMMPS_NS::LAMMPS::git_branch() { return "(unknown)"; }\n'
                     'const char *LAMMPS_NS::LAMMPS::git_descriptor() { return "(unknown)"; }\n')


def main():
    if not os.path.isdir(SRC):
        print("build_ref: %s absent -- keeping prebuilt oracle/_ref as is" % REF)
        return 0
    os.makedirs(OBJ, exist_ok=True)
    srcs = selected_sources()
    gen_headers(srcs)
    inc = ["-I" + GEN, "-I" + SRC, "-I" + os.path.join(SRC, "STUBS"),
           "-I" + os.path.join(SRC, "MANYBODY"), "-I" + os.path.join(SRC, "MOLECULE"),
           "-I" + os.path.join(SRC, "OPENMP")]
    flags = ["-O3", "-DNDEBUG", "-ffp-contract=off", "-fopenmp", "-fPIC", "-std=c++11",
             "-DLAMMPS_SMALLBIG", "-DLAMMPS_EXCEPTIONS", "-DLMP_OPENMP", "-w"]
    nj = ["cxx = %s" % CXX, "flags = %s %s" % (" ".join(flags), " ".join(inc)),
          "rule cc", "  command = $cxx $flags -MMD -MF $out.d -c $in -o $out",
          "  depfile = $out.d", "  deps = gcc", "  description = CC $in",
          "rule so", "  command = $cxx -shared -fopenmp -o $out $in", "  description = LINK $out",
          "rule exe", "  command = $cxx -fopenmp -o $out $in -Wl,-rpath,'$$ORIGIN' -L%s -llammps_ref" % OUT,
          "  description = LINK $out",
          ""]
    objs = []
    main_obj = None
    for s in srcs:
        tag = os.path.relpath(s, SRC).replace("/", "_")[:-4] + ".o"
        o = os.path.join(OBJ, tag)
        nj.append("build %s: cc %s" % (o, s))
        if os.path.basename(s) == "main.cpp" and os.path.dirname(s) == SRC:
            main_obj = o
        else:
            objs.append(o)
    lib = os.path.join(OUT, "liblammps_ref.so")
    nj.append("build %s: so %s" % (lib, " ".join(objs)))
    nj.append("build %s: exe %s | %s" % (os.path.join(OUT, "lmp"), main_obj, lib))
    nj.append("default %s" % os.path.join(OUT, "lmp"))
    write_if_changed(os.path.join(OUT, "build.ninja"), "\n".join(nj) + "\n")
    r = subprocess.run(["ninja", "-C", OUT, "-j", str(os.cpu_count() or 4)])
    if r.returncode != 0:
        return r.returncode
    pot = os.path.join(OUT, "potentials")
    os.makedirs(pot, exist_ok=True)
    for src in [os.path.join(REF, "bench", "Cu_u3.eam"),
                os.path.join(REF, "potentials", "Cu_mishin1.eam.alloy"),
                os.path.join(REF, "potentials", "Al_jnp.eam"),
                os.path.join(REF, "potentials", "CuNi.eam.alloy"),
                os.path.join(REF, "potentials", "AlFe_mm.eam.fs")]:
        dst = os.path.join(pot, os.path.basename(src))
        if os.path.exists(src) and not os.path.exists(dst):
            shutil.copyfile(src, dst)
    data = os.path.join(OUT, "data")
    os.makedirs(data, exist_ok=True)
    if not os.path.exists(os.path.join(data, "data.chain")):
        shutil.copyfile(os.path.join(REF, "bench", "data.chain"), os.path.join(data, "data.chain"))
    print("build_ref: ok ->", OUT)
    return 0


if __name__ == "__main__":
    sys.exit(main())
