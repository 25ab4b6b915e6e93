"""TEST INFRASTRUCTURE ONLY — ctypes wrappers of the two checkers:

  Oracle      oracle/libmd_oracle.so   the plain-C restatement (oracle/md_oracle.c)
  Reference   oracle/_ref/liblammps_ref.so   the UNMODIFIED reference built by oracle/build_ref.py,
              driven through its own C library API (/root/reference/src/library.h:123-296)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product path (lammps-kokkos-port_b200/) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(HERE, "libmd_oracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_SO = os.path.join(REF_DIR, "liblammps_ref.so")
REF_EXE = os.path.join(REF_DIR, "lmp")
POT_DIR = os.path.join(REF_DIR, "potentials")


def build_oracle():
    subprocess.run(["make", "-C", HERE, "CC=gcc"], check=True, stdout=subprocess.DEVNULL)


class OraThermo(C.Structure):
    _fields_ = [("step", C.c_long), ("natoms", C.c_double), ("temp", C.c_double), ("pe", C.c_double),
                ("ke", C.c_double), ("press", C.c_double), ("virial", C.c_double * 6), ("emol", C.c_double)]

    def as_dict(self):
        return dict(step=self.step, natoms=self.natoms, temp=self.temp, pe=self.pe, ke=self.ke, press=self.press,
                    virial=list(self.virial), emol=self.emol)


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Oracle:
    def __init__(self, boxlo, boxhi, x, v, type_, tag, mass, units):
        if not os.path.exists(ORACLE_SO):
            build_oracle()
        L = self.L = C.CDLL(ORACLE_SO)
        L.ora_create.restype = C.c_void_p
        for fn in ("ora_x", "ora_v", "ora_f", "ora_tag", "ora_type", "ora_image", "ora_numneigh", "ora_eatom", "ora_vatom"):
            getattr(L, fn).restype = C.c_void_p
        L.ora_neigh_total.restype = C.c_long
        x = _d(x); self.n = x.shape[0]
        v = _d(v); type_ = _i(type_); tag = _i(tag); mass = _d(mass)
        self.ntypes = mass.shape[0] - 1
        lo = (C.c_double * 3)(*boxlo); hi = (C.c_double * 3)(*boxhi)
        self.h = C.c_void_p(L.ora_create(lo, hi, self.n, _p(x), _p(v), _p(type_), _p(tag), self.ntypes, _p(mass),
                                         C.c_double(units["boltz"]), C.c_double(units["mvv2e"]),
                                         C.c_double(units["ftm2v"]), C.c_double(units["nktv2p"])))

    def close(self):
        if self.h:
            self.L.ora_destroy(self.h)
            self.h = None

    def velocity_finalize(self, t):
        self.L.ora_velocity_finalize(self.h, C.c_double(t))

    def neigh_config(self, cutneighsq, cutneighmax, skin, every, delay, check, half):
        cn = _d(cutneighsq)
        self.L.ora_neigh_config(self.h, _p(cn), C.c_double(cutneighmax), C.c_double(skin), every, delay, check, int(half))

    def pair_lj(self, T):
        a = [_d(T[k]) for k in ("cutsq", "lj1", "lj2", "lj3", "lj4", "offset")]
        self.L.ora_pair_lj(self.h, *[_p(t) for t in a])

    def pair_eam(self, tb):
        fr, rr, zr = _d(tb["frho_spline"]), _d(tb["rhor_spline"]), _d(tb["z2r_spline"])
        t2f, t2r, t2z, sc = _i(tb["type2frho"]), _i(tb["type2rhor"]), _i(tb["type2z2r"]), _d(tb["scale"])
        self.L.ora_pair_eam(self.h, int(tb["nrho"]), int(tb["nr"]), C.c_double(tb["rdr"]), C.c_double(tb["rdrho"]),
                            C.c_double(tb["rhomax"]), C.c_double(tb["cutforcesq"]), fr.shape[0], _p(fr), rr.shape[0],
                            _p(rr), zr.shape[0], _p(zr), _p(t2f), _p(t2r), _p(t2z), _p(sc))

    def set_topology(self, tp):
        a = {k: _i(tp[k]) for k in ("bond_offset", "bond_partner", "bond_type", "special_offset", "special_tag",
                                     "special_level")}
        slj, sco = _d(tp["special_lj"]), _d(tp["special_coul"])
        if self.L.ora_set_topology(self.h, int(tp["natoms"]), _p(a["bond_offset"]), _p(a["bond_partner"]),
                                   _p(a["bond_type"]), _p(a["special_offset"]), _p(a["special_tag"]),
                                   _p(a["special_level"]), _p(slj), _p(sco)):
            raise RuntimeError("oracle: special_bonds weights other than 0 or 1 are not restated")

    def bond_fene(self, k, r0, epsilon, sigma):
        a = [_d(t) for t in (k, r0, epsilon, sigma)]
        self.L.ora_bond_fene(self.h, len(a[0]) - 1, *[_p(t) for t in a])

    def atom_sort(self):
        """Atom::sort as Verlet::setup calls it (after neigh_config, before setup)"""
        self.L.ora_atom_sort(self.h)

    def fix_langevin(self, t_start, t_stop, t_period, seed, dt, ratio=None):
        r = _d(ratio) if ratio is not None else None
        if self.L.ora_fix_langevin(self.h, C.c_double(t_start), C.c_double(t_stop), C.c_double(t_period), int(seed),
                                   _p(r) if r is not None else None, C.c_double(dt)):
            raise RuntimeError("oracle: fix langevin needs period > 0 and 0 < seed <= 900000000")

    def neigh_build(self):
        if self.L.ora_neigh_build(self.h):
            raise RuntimeError("oracle: neighbor build failed (too many ghost swaps)")

    def compute(self, eflag=1, vflag=1):
        if self.L.ora_compute(self.h, eflag, vflag):
            raise RuntimeError("oracle: compute needs the half list and a pair style")

    def peratom(self, eflag_atom=1, vflag_atom=1):
        """(eatom[nlocal], vatom[nlocal,6]) pair + bond, local order (see atoms()["tag"]); vatom is the
        raw virial sum: `compute stress/atom NULL virial` prints -nktv2p * vatom"""
        if self.L.ora_compute_peratom(self.h, eflag_atom, vflag_atom):
            raise RuntimeError("oracle: per-atom compute failed")
        n = self.L.ora_nlocal(self.h)
        return self._arr("ora_eatom", (n,), np.float64), self._arr("ora_vatom", (n, 6), np.float64)

    def setup(self):
        if self.L.ora_setup(self.h):
            raise RuntimeError("oracle: setup failed")

    def thermo(self):
        t = OraThermo()
        self.L.ora_thermo_compute(self.h, C.byref(t))
        return t.as_dict()

    def run(self, nsteps, dt, thermo_every=0):
        cap = (nsteps // thermo_every + 2) if thermo_every > 0 else 2
        arr = (OraThermo * cap)()
        n = self.L.ora_run(self.h, nsteps, C.c_double(dt), thermo_every, arr, cap)
        if n < 0:
            raise RuntimeError("oracle: run failed")
        return [arr[k].as_dict() for k in range(n)]

    def _arr(self, fn, shape, dtype):
        ptr = getattr(self.L, fn)(self.h)
        n = int(np.prod(shape))
        ct = C.c_double if dtype == np.float64 else C.c_int
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(ct)), shape=(n,)).reshape(shape).copy()

    def nlocal(self):
        return self.L.ora_nlocal(self.h), self.L.ora_nghost(self.h)

    def nbuilds(self):
        return self.L.ora_nbuilds(self.h)

    def atoms(self):
        n = self.L.ora_nlocal(self.h)
        return dict(x=self._arr("ora_x", (n, 3), np.float64), v=self._arr("ora_v", (n, 3), np.float64),
                    f=self._arr("ora_f", (n, 3), np.float64), tag=self._arr("ora_tag", (n,), np.int32),
                    type=self._arr("ora_type", (n,), np.int32), image=self._arr("ora_image", (n,), np.int32))

    def ghosts(self):
        n, g = self.nlocal()
        x = self._arr("ora_x", (n + g, 3), np.float64)[n:]
        tag = self._arr("ora_tag", (n + g,), np.int32)[n:]
        return dict(x=x, tag=tag)

    def neigh_tags(self):
        n = self.L.ora_nlocal(self.h)
        nn = self._arr("ora_numneigh", (n,), np.int32)
        tot = self.L.ora_neigh_total(self.h)
        flat = np.empty(tot, np.int32)
        self.L.ora_neigh_tags(self.h, _p(flat))
        return nn, flat

    def bin_geometry(self):
        a = (C.c_int * 12)(); b = (C.c_double * 6)()
        self.L.ora_bin_geometry(self.h, a, b)
        return list(a), list(b)


def ranmars_stream(seed, n):
    """first n RanMars::uniform() draws after construction (src/random_mars.cpp), from the C oracle"""
    if not os.path.exists(ORACLE_SO):
        build_oracle()
    L = C.CDLL(ORACLE_SO)
    out = np.empty(n, np.float64)
    if L.ora_ranmars_stream(int(seed), C.c_long(n), _p(out)):
        raise RuntimeError("Invalid seed for Marsaglia random # generator")
    return out


class Reference:
    """The unmodified reference through its C library API (src/library.h)."""

    def __init__(self, args=("-log", "none", "-screen", "none", "-echo", "none", "-nocite")):
        if not os.path.exists(REF_SO):
            raise RuntimeError("oracle/_ref/liblammps_ref.so missing: run oracle/build_ref.py where /root/reference exists")
        L = self.L = C.CDLL(REF_SO, mode=C.RTLD_GLOBAL)
        L.lammps_open_no_mpi.restype = C.c_void_p
        L.lammps_extract_atom.restype = C.c_void_p
        L.lammps_extract_global.restype = C.c_void_p
        L.lammps_get_thermo.restype = C.c_double
        L.lammps_get_natoms.restype = C.c_double
        argv = [b"lammps"] + [a.encode() for a in args]
        arr = (C.c_char_p * len(argv))(*argv)
        self.h = C.c_void_p(L.lammps_open_no_mpi(len(argv), arr, None))

    def close(self):
        if self.h:
            self.L.lammps_close(self.h)
            self.h = None

    def command(self, cmd):
        for line in cmd.strip().split("\n"):
            line = line.strip()
            if line:
                self.L.lammps_command(self.h, line.encode())
                if self.L.lammps_has_error(self.h):
                    buf = C.create_string_buffer(512)
                    self.L.lammps_get_last_error_message(self.h, buf, 512)
                    raise RuntimeError("reference: " + buf.value.decode())

    def natoms(self):
        return int(self.L.lammps_get_natoms(self.h))

    def setting(self, name):
        return self.L.lammps_extract_setting(self.h, name.encode())

    def thermo(self, kw):
        return self.L.lammps_get_thermo(self.h, kw.encode())

    def gather(self, name, count, dtype=np.float64):
        n = self.natoms()
        out = np.empty(n * count, dtype=dtype)
        self.L.lammps_gather_atoms(self.h, name.encode(), 1 if dtype == np.float64 else 0, count, _p(out))
        return out.reshape(n, count) if count > 1 else out

    def local_ids(self):
        """tags of owned + ghost atoms in local order"""
        nall = self.setting("nlocal") + self.setting("nghost")
        ptr = self.L.lammps_extract_atom(self.h, b"id")
        return np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_int)), shape=(nall,)).copy()

    def local_x(self):
        nall = self.setting("nlocal") + self.setting("nghost")
        ptr = self.L.lammps_extract_atom(self.h, b"x")  # double**, rows contiguous (src/memory.h:144-160)
        row0 = C.cast(ptr, C.POINTER(C.POINTER(C.c_double)))[0]
        return np.ctypeslib.as_array(row0, shape=(nall * 3,)).reshape(nall, 3).copy()

    def _list_by_tag(self, idx):
        ids = self.local_ids()
        n = self.L.lammps_neighlist_num_elements(self.h, idx)
        out = {}
        iatom = C.c_int(); nn = C.c_int(); ptr = C.POINTER(C.c_int)()
        for e in range(n):
            self.L.lammps_neighlist_element_neighbors(self.h, idx, e, C.byref(iatom), C.byref(nn), C.byref(ptr))
            js = np.ctypeslib.as_array(ptr, shape=(nn.value,)) & 0x1FFFFFFF if nn.value else np.empty(0, np.int32)
            out[int(ids[iatom.value])] = ids[js].copy()
        return out

    def neighlist_by_tag(self, style):
        """perpetual (half, newton on) pair list of `style` as {tag_i: [tag_j...]} (library.h:219-223)"""
        idx = self.L.lammps_find_pair_neighlist(self.h, style.encode(), 1, 0, 0)
        if idx < 0:
            raise RuntimeError("reference: no neighbor list for pair style " + style)
        return self._list_by_tag(idx)

    def full_neighlist_by_tag(self, compute_id):
        """the full list (NPairFullBinAtomonly) requested by a `compute coord/atom` with that id"""
        idx = self.L.lammps_find_compute_neighlist(self.h, compute_id.encode(), 0)
        if idx < 0:
            raise RuntimeError("reference: no neighbor list for compute " + compute_id)
        return self._list_by_tag(idx)


IN_LJ = """
units lj
atom_style atomic
atom_modify map array
lattice fcc 0.8442
region box block 0 {nx} 0 {ny} 0 {nz}
create_box 1 box
create_atoms 1 box
mass 1 1.0
velocity all create 1.44 87287 loop geom
pair_style lj/cut 2.5
pair_coeff 1 1 1.0 1.0 2.5
neighbor 0.3 bin
neigh_modify delay 0 every 20 check no
fix 1 all nve
"""

IN_EAM = """
units metal
atom_style atomic
atom_modify map array
lattice fcc 3.615
region box block 0 {nx} 0 {ny} 0 {nz}
create_box 1 box
create_atoms 1 box
pair_style {style}
pair_coeff {coeff}
velocity all create 1600.0 376847 loop geom
neighbor 1.0 bin
neigh_modify every 1 delay 5 check yes
fix 1 all nve
timestep 0.005
"""


def ref_exe_run(script_text, threads=1, omp=False, timeout=600):
    """Run the reference executable on an input script; return (loop_time_s, atoms, steps, stdout)."""
    import re
    import tempfile
    with tempfile.NamedTemporaryFile("w", suffix=".in", delete=False) as fh:
        fh.write(script_text)
        path = fh.name
    cmd = [REF_EXE, "-in", path, "-log", "none", "-echo", "none", "-nocite"]
    if omp:
        cmd += ["-sf", "omp", "-pk", "omp", str(threads)]
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, env=env)
    os.unlink(path)
    m = re.search(r"Loop time of ([0-9.eE+-]+) on \d+ procs for (\d+) steps with (\d+) atoms", r.stdout)
    if not m:
        raise RuntimeError("reference run failed:\n" + r.stdout[-2000:] + r.stderr[-2000:])
    return float(m.group(1)), int(m.group(3)), int(m.group(2)), r.stdout
