"""ctypes binding of libb200md.so (include/b200md.h) plus a thin host-side driver that mirrors
the reference's input-script vocabulary for the hot path (units / lattice / region /
create_atoms / mass / velocity / pair_style / pair_coeff / neighbor / neigh_modify / fix nve /
run), i.e. what `bench/in.lj` and `bench/in.eam` (/root/reference/bench) set up.

PyTorch is not needed here; torch.distributed is used by callers only to ship the NCCL
unique id between ranks.  There is no CPU fallback: every compute call goes to the CUDA
library and raises B200MDError (with the library's message) when it fails.
"""
import ctypes as C
import math
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200MD_LIB", os.path.join(_HERE, "libb200md.so"))

LIST_FULL, LIST_HALF_NEWTON = 0, 1
PERATOM_PAIR, PERATOM_BOND = 1, 2
MAXSWAP = 64
OPTION_KEYS = ("halo", "overlap", "split", "fuse_nve", "eam_smem", "eam_mask", "bin_scale", "cell_order", "self_border", "phase_timing", "lanes", "build_lanes", "graph", "cluster", "trace")


class B200MDError(RuntimeError):
    pass


class SwapSchedule(C.Structure):
    _fields_ = [("nswap", C.c_int), ("maxneed", C.c_int * 3), ("dim", C.c_int * MAXSWAP),
                ("ineed", C.c_int * MAXSWAP), ("sendflag", C.c_int * MAXSWAP),
                ("sendproc", C.c_int * MAXSWAP), ("recvproc", C.c_int * MAXSWAP),
                ("pbc", (C.c_int * 3) * MAXSWAP), ("slablo", C.c_double * MAXSWAP),
                ("slabhi", C.c_double * MAXSWAP)]


class BinGeometry(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("nbinx", "nbiny", "nbinz", "mbinx", "mbiny", "mbinz", "mbinxlo",
                                        "mbinylo", "mbinzlo", "sx", "sy", "sz")] + \
               [(n, C.c_double) for n in ("binsizex", "binsizey", "binsizez", "bininvx", "bininvy", "bininvz")]


class Thermo(C.Structure):
    _fields_ = [("step", C.c_longlong), ("natoms", C.c_double), ("temp", C.c_double), ("pe", C.c_double),
                ("ke", C.c_double), ("press", C.c_double), ("virial", C.c_double * 6), ("emol", C.c_double)]

    def as_dict(self):
        return dict(step=self.step, natoms=self.natoms, temp=self.temp, pe=self.pe, ke=self.ke,
                    press=self.press, virial=list(self.virial), emol=self.emol)


class Timing(C.Structure):
    _fields_ = [("total_ms", C.c_double), ("pair_ms", C.c_double), ("neigh_ms", C.c_double),
                ("comm_ms", C.c_double), ("modify_ms", C.c_double), ("launches", C.c_longlong),
                ("pair_launches", C.c_longlong), ("nbuilds", C.c_int)]


_lib = None


def lib():
    """Load libb200md.so (built in-tree by __graft_entry__.build()); fail loudly if absent."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise B200MDError("libb200md.so is not built (%s); run __graft_entry__.build()" % LIB_PATH)
        # bind the same NCCL torch ships (site-packages/nvidia/nccl) so that load order of torch and
        # this library does not matter; falls back to the system libnccl.so.2
        try:
            import importlib.util
            spec = importlib.util.find_spec("nvidia.nccl")
            if spec and spec.submodule_search_locations:
                C.CDLL(os.path.join(list(spec.submodule_search_locations)[0], "lib", "libnccl.so.2"), mode=C.RTLD_GLOBAL)
        except OSError:
            pass
        _lib = C.CDLL(LIB_PATH)
        _lib.b200md_last_error.restype = C.c_char_p
        _lib.b200md_version.restype = C.c_char_p
    return _lib


def _chk(rc):
    if rc != 0:
        raise B200MDError(lib().b200md_last_error().decode())


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def _d3(v):
    return (C.c_double * 3)(*[float(x) for x in v])


def _i3(v):
    return (C.c_int * 3)(*[int(x) for x in v])


# ---- host-only helpers -------------------------------------------------------------------
def procgrid(nranks, prd):
    out = (C.c_int * 3)()
    _chk(lib().b200md_procgrid(int(nranks), _d3(prd), out))
    return list(out)


def rank2loc(rank, grid):
    out = (C.c_int * 3)()
    _chk(lib().b200md_rank2loc(int(rank), _i3(grid), out))
    return list(out)


def local_box(boxlo, boxhi, grid, myloc):
    lo, hi = (C.c_double * 3)(), (C.c_double * 3)()
    _chk(lib().b200md_local_box(_d3(boxlo), _d3(boxhi), _i3(grid), _i3(myloc), lo, hi))
    return list(lo), list(hi)


def comm_schedule(boxlo, boxhi, periodic, grid, myloc, cutghost):
    s = SwapSchedule()
    _chk(lib().b200md_comm_schedule(_d3(boxlo), _d3(boxhi), _i3(periodic), _i3(grid), _i3(myloc),
                                    C.c_double(cutghost), C.byref(s)))
    return s


def halo_directions(boxlo, boxhi, periodic, grid, myloc, cutghost):
    """direct-halo tables (csrc/halo26.cu): peer[27], shift[27][3], slabhi_left[3], slablo_right[3], usable"""
    peer = (C.c_int * 27)(); shift = ((C.c_int * 3) * 27)()
    shl, slr = (C.c_double * 3)(), (C.c_double * 3)()
    ok = C.c_int()
    _chk(lib().b200md_halo_directions(_d3(boxlo), _d3(boxhi), _i3(periodic), _i3(grid), _i3(myloc), C.c_double(cutghost),
                                      peer, shift, shl, slr, C.byref(ok)))
    return list(peer), [list(r) for r in shift], list(shl), list(slr), bool(ok.value)


def atom_sort_order(x, sublo, subhi, cutneighmax):
    """Atom::sort's permutation (src/atom.cpp:2246-2349): new atom I = old atom perm[I]"""
    x = _d(x)
    perm = np.empty(x.shape[0], np.int32)
    _chk(lib().b200md_atom_sort_order(int(x.shape[0]), _p(x), _d3(sublo), _d3(subhi), C.c_double(cutneighmax), _p(perm)))
    return perm


def ranmars_state(seed):
    """RanMars(seed) after construction, integer form (src/random_mars.cpp:30-72): uint32[97]"""
    st = np.zeros(97, np.uint32)
    _chk(lib().b200md_ranmars_state(int(seed), _p(st)))
    return st


def ranmars_jump(state, k):
    st = np.ascontiguousarray(state, np.uint32)
    out = np.zeros(97, np.uint32)
    _chk(lib().b200md_ranmars_jump(_p(st), C.c_longlong(int(k)), _p(out)))
    return out


def ranmars_draws(state, position, count):
    st = np.ascontiguousarray(state, np.uint32)
    out = np.empty(count, np.float64)
    _chk(lib().b200md_ranmars_draws(_p(st), C.c_longlong(int(position)), int(count), _p(out)))
    return out


def bin_setup(boxlo, boxhi, sublo, subhi, cutghost, cutneighmax):
    g = BinGeometry()
    _chk(lib().b200md_bin_setup(_d3(boxlo), _d3(boxhi), _d3(sublo), _d3(subhi), C.c_double(cutghost),
                                C.c_double(cutneighmax), C.byref(g)))
    return g


def lattice_fcc(scale, nx, ny, nz, sublo, subhi, seed, mass):
    """Atoms of `lattice fcc` + `create_atoms 1 box` inside [sublo,subhi) with the raw
    `velocity ... loop geom` draws (host.cpp)."""
    n = C.c_int(0)
    L = lib()
    _chk(L.b200md_lattice_fcc(C.c_double(scale), nx, ny, nz, _d3(sublo), _d3(subhi), int(seed), C.c_double(mass),
                              C.byref(n), None, None, None))
    x = np.empty((n.value, 3)); v = np.empty((n.value, 3)); tag = np.empty(n.value, dtype=np.int32)
    _chk(L.b200md_lattice_fcc(C.c_double(scale), nx, ny, nz, _d3(sublo), _d3(subhi), int(seed), C.c_double(mass),
                              C.byref(n), _p(x), _p(v), _p(tag)))
    return x, v, tag


def pinned_empty(shape, dtype=np.float64):
    """numpy array over cudaMallocHost memory (b200md_host_alloc); kept alive by the array's base."""
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    ptr = C.c_void_p()
    _chk(lib().b200md_host_alloc(C.byref(ptr), C.c_size_t(max(n, 8))))
    buf = (C.c_char * max(n, 8)).from_address(ptr.value)
    return np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)


KE_STATIC_MAX_ATOMS = 16777216     # B200MD_KE_STATIC_MAX_ATOMS (include/b200md.h)


def ke_sums_static(n, stride, v_soa_dev, type_dev, mass_dev, out5_dev, stream=None):
    """b200md_ke_sums_static: {sum m v.v, sum m vx, sum m vy, sum m vz, sum m} of n atoms into out5_dev, one launch on
    `stream`, nothing allocated (scratch in a static device array of the library).  Every *_dev argument is a DEVICE
    address (int) on the current device; type_dev may be None/0.  Not synchronised.  No counterpart in the reference."""
    L = lib()
    L.b200md_ke_sums_static.argtypes = [C.c_int, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    _chk(L.b200md_ke_sums_static(int(n), int(stride), v_soa_dev or None, type_dev or None, mass_dev or None,
                                 out5_dev or None, stream or None))


# ---- unit systems (src/update.cpp:145-200) ---------------------------------------------------
UNITS = {
    "lj": dict(boltz=1.0, mvv2e=1.0, ftm2v=1.0, nktv2p=1.0, dt=0.005, skin=0.3),
    "metal": dict(boltz=8.617343e-5, mvv2e=1.0364269e-4, ftm2v=1.0 / 1.0364269e-4, nktv2p=1.6021765e6,
                  dt=0.001, skin=2.0),
    "real": dict(boltz=0.0019872067, mvv2e=48.88821291 * 48.88821291, ftm2v=1.0 / 48.88821291 / 48.88821291,
                 nktv2p=68568.415, dt=1.0, skin=2.0),
}


class Context:
    """One b200md_ctx (one GPU, one rank)."""

    def __init__(self, device=0, rank=0, nranks=1):
        self.L = lib()
        self.h = C.c_void_p()
        _chk(self.L.b200md_create(int(device), int(rank), int(nranks), C.byref(self.h)))
        self.rank, self.nranks = rank, nranks
        self.ntypes = 0
        # developer convenience of THIS driver (the library itself reads no environment): B200MD_<KEY>=value
        for key in OPTION_KEYS:
            v = os.environ.get("B200MD_" + key.upper())
            if v is not None:
                self.set_option(key, v)

    def set_option(self, key, value):
        """`package b200 <key> <value>` (include/b200md.h: b200md_set_option)"""
        if isinstance(value, bool):
            value = "on" if value else "off"
        _chk(self.L.b200md_set_option(self.h, str(key).encode(), str(value).encode()))

    def get_option(self, key):
        buf = C.create_string_buffer(64)
        _chk(self.L.b200md_get_option(self.h, str(key).encode(), buf, 64))
        return buf.value.decode()

    def close(self):
        if self.h:
            self.L.b200md_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # comm
    @staticmethod
    def unique_id():
        buf = C.create_string_buffer(128)
        _chk(lib().b200md_comm_unique_id(buf))
        return buf.raw

    def comm_init(self, id128):
        _chk(self.L.b200md_comm_init(self.h, C.c_char_p(id128)))

    def set_units(self, name):
        u = UNITS[name]
        _chk(self.L.b200md_set_units(self.h, C.c_double(u["boltz"]), C.c_double(u["mvv2e"]), C.c_double(u["ftm2v"]),
                                     C.c_double(u["nktv2p"])))
        self.units = u

    def set_box(self, boxlo, boxhi, periodic=(1, 1, 1), grid=(1, 1, 1)):
        _chk(self.L.b200md_set_box(self.h, _d3(boxlo), _d3(boxhi), _i3(periodic), _i3(grid)))

    def subbox(self):
        lo, hi = (C.c_double * 3)(), (C.c_double * 3)()
        _chk(self.L.b200md_get_subbox(self.h, lo, hi))
        return list(lo), list(hi)

    def set_atoms(self, x, v, type_, tag, mass, image=None):
        x = _d(x); n = x.shape[0]
        v = _d(v) if v is not None else None
        type_ = _i(type_); tag = _i(tag)
        mass = _d(mass)
        self.ntypes = mass.shape[0] - 1
        self.mass = mass.copy()
        image = _i(image) if image is not None else None
        _chk(self.L.b200md_set_atoms(self.h, n, _p(x), _p(v), _p(type_), _p(tag), _p(image), self.ntypes, _p(mass)))

    def nlocal(self):
        a, b = C.c_int(), C.c_int()
        _chk(self.L.b200md_get_nlocal(self.h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def get_atoms(self, fields=("x", "v", "f", "type", "tag", "image")):
        n, _ = self.nlocal()
        out = {}
        arr = dict(x=np.empty((n, 3)), v=np.empty((n, 3)), f=np.empty((n, 3)), type=np.empty(n, np.int32),
                   tag=np.empty(n, np.int32), image=np.empty(n, np.int32))
        g = lambda k: _p(arr[k]) if k in fields else None
        _chk(self.L.b200md_get_atoms(self.h, g("x"), g("v"), g("f"), g("type"), g("tag"), g("image")))
        for k in fields:
            out[k] = arr[k]
        return out

    def get_atoms_into(self, x=None, v=None, f=None, tag=None):
        """download into caller-provided (pinned) arrays; no allocation"""
        g = lambda a: _p(a) if a is not None else None
        _chk(self.L.b200md_get_atoms(self.h, g(x), g(v), g(f), None, g(tag), None))

    def get_atoms_async(self, x=None, f=None):
        """enqueue the download of x / f into pinned arrays (copy stream); wait_download(False) = x is there"""
        g = lambda a: _p(a) if a is not None else None
        _chk(self.L.b200md_get_atoms_async(self.h, g(x), g(f)))

    def wait_download(self, everything=True):
        _chk(self.L.b200md_wait_download(self.h, int(everything)))

    def wait_download_part(self, part):
        """piece `part` of the x download of get_atoms_async is on the host: (first, count), count 0 past the last piece"""
        first, count = C.c_int(), C.c_int()
        _chk(self.L.b200md_wait_download_part(self.h, int(part), C.byref(first), C.byref(count)))
        return first.value, count.value

    def put_atoms_range(self, x, first, count):
        """upload the positions of the owned atoms [first, first+count) from the host array x (row 0 = atom 0)"""
        _chk(self.L.b200md_put_atoms_range(self.h, _p(x), int(first), int(count)))

    def put_atoms(self, x=None, v=None):
        g = lambda a: _p(a) if a is not None else None
        _chk(self.L.b200md_put_atoms(self.h, g(x), g(v)))

    def get_ghosts(self):
        _, n = self.nlocal()
        x = np.empty((n, 3)); t = np.empty(n, np.int32); tag = np.empty(n, np.int32)
        _chk(self.L.b200md_get_ghosts(self.h, _p(x), _p(t), _p(tag)))
        return dict(x=x, type=t, tag=tag)

    def velocity_finalize(self, t_desired, zero_momentum=True):
        _chk(self.L.b200md_velocity_finalize(self.h, C.c_double(t_desired), int(zero_momentum)))

    # neighbor
    def neigh_config(self, cutneighsq, cutneighmax, skin, every=1, delay=0, check=1, list_kind=LIST_FULL):
        cn = _d(cutneighsq)
        _chk(self.L.b200md_neigh_config(self.h, self.ntypes, _p(cn), C.c_double(cutneighmax), C.c_double(skin),
                                        int(every), int(delay), int(check), int(list_kind)))

    def neigh_decide(self):
        r = C.c_int()
        _chk(self.L.b200md_neigh_decide(self.h, C.byref(r)))
        return r.value

    def neigh_build(self):
        _chk(self.L.b200md_neigh_build(self.h))

    def forward_comm(self):
        _chk(self.L.b200md_forward_comm(self.h))

    def reverse_comm(self):
        _chk(self.L.b200md_reverse_comm(self.h))

    def neigh_export(self, by_tag=True):
        n, _ = self.nlocal()
        nn = np.empty(n, np.int32)
        tot = C.c_longlong()
        _chk(self.L.b200md_neigh_export(self.h, int(by_tag), _p(nn), None, C.byref(tot)))
        flat = np.empty(tot.value, np.int32)
        _chk(self.L.b200md_neigh_export(self.h, int(by_tag), _p(nn), _p(flat), C.byref(tot)))
        return nn, flat

    def neigh_stats(self):
        a, b, t, p = C.c_longlong(), C.c_longlong(), C.c_longlong(), C.c_int()
        _chk(self.L.b200md_neigh_stats(self.h, C.byref(a), C.byref(b), C.byref(t), C.byref(p)))
        return dict(nbuilds=a.value, ndanger=b.value, total=t.value, pitch=p.value)

    def neigh_layout(self):
        cl, ent, lanes = C.c_int(), C.c_longlong(), C.c_int()
        _chk(self.L.b200md_neigh_layout(self.h, C.byref(cl), C.byref(ent), C.byref(lanes)))
        return dict(cluster=bool(cl.value), entries=ent.value, lanes=lanes.value)

    def set_lanes_per_atom(self, n):
        _chk(self.L.b200md_set_lanes_per_atom(self.h, int(n)))

    # pair
    def pair_lj_cut_config(self, cutsq, lj1, lj2, lj3, lj4, offset):
        a = [_d(t) for t in (cutsq, lj1, lj2, lj3, lj4, offset)]
        _chk(self.L.b200md_pair_lj_cut_config(self.h, self.ntypes, *[_p(t) for t in a]))

    def pair_eam_config(self, tb):
        """tb: dict with the PairEAM spline tables (see eam_host.py)."""
        fr, rr, zr = _d(tb["frho_spline"]), _d(tb["rhor_spline"]), _d(tb["z2r_spline"])
        t2f, t2r, t2z = _i(tb["type2frho"]), _i(tb["type2rhor"]), _i(tb["type2z2r"])
        sc = _d(tb["scale"])
        _chk(self.L.b200md_pair_eam_config(self.h, self.ntypes, int(tb["nrho"]), int(tb["nr"]), C.c_double(tb["rdr"]),
                                           C.c_double(tb["rdrho"]), C.c_double(tb["rhomax"]),
                                           C.c_double(tb["cutforcesq"]), fr.shape[0], _p(fr), rr.shape[0], _p(rr),
                                           zr.shape[0], _p(zr), _p(t2f), _p(t2r), _p(t2z), _p(sc)))

    def pair_compute(self, eflag=1, vflag=1):
        e = C.c_double(); vir = (C.c_double * 6)()
        _chk(self.L.b200md_pair_compute(self.h, int(eflag), int(vflag), C.byref(e), vir))
        return e.value, list(vir)

    # molecular topology / bonds
    def set_topology(self, tp):
        """tp: dict from molecule_host.build_topology"""
        a = {k: _i(tp[k]) for k in ("bond_offset", "bond_partner", "bond_type", "special_offset", "special_tag",
                                     "special_level")}
        slj = (C.c_double * 4)(*[float(t) for t in tp["special_lj"]])
        sco = (C.c_double * 4)(*[float(t) for t in tp["special_coul"]])
        _chk(self.L.b200md_set_topology(self.h, C.c_longlong(int(tp["natoms"])), _p(a["bond_offset"]),
                                        _p(a["bond_partner"]), _p(a["bond_type"]), _p(a["special_offset"]),
                                        _p(a["special_tag"]), _p(a["special_level"]), slj, sco))

    def bond_fene_config(self, k, r0, epsilon, sigma):
        """1-based per-bond-type arrays (index 0 unused)"""
        a = [_d(t) for t in (k, r0, epsilon, sigma)]
        _chk(self.L.b200md_bond_fene_config(self.h, len(a[0]) - 1, *[_p(t) for t in a]))

    def bond_compute(self, eflag=1, vflag=1):
        e = C.c_double(); vir = (C.c_double * 6)()
        _chk(self.L.b200md_bond_compute(self.h, int(eflag), int(vflag), C.byref(e), vir))
        return e.value, list(vir)

    def bond_tallies(self):
        e = C.c_double(); vir = (C.c_double * 6)()
        _chk(self.L.b200md_bond_tallies(self.h, C.byref(e), vir))
        return e.value, list(vir)

    def pair_tallies(self):
        e = C.c_double(); vir = (C.c_double * 6)()
        _chk(self.L.b200md_pair_tallies(self.h, C.byref(e), vir))
        return e.value, list(vir)

    def peratom(self, eflag_atom=1, vflag_atom=1, which=PERATOM_PAIR | PERATOM_BOND):
        """(eatom[nlocal], vatom[nlocal,6]) of the current configuration in get_atoms() order: what
        `compute pe/atom` holds, and `compute stress/atom` before its kinetic term and -nktv2p factor
        (src/compute_pe_atom.cpp:92-170, src/compute_stress_atom.cpp:140-240); None for a part not asked for"""
        n, _ = self.nlocal()
        e = np.zeros(n) if eflag_atom else None
        v = np.zeros((n, 6)) if vflag_atom else None
        _chk(self.L.b200md_peratom(self.h, int(eflag_atom), int(vflag_atom), int(which),
                                   _p(e) if eflag_atom else None, _p(v) if vflag_atom else None))
        return e, v

    def fix_langevin(self, t_start, t_stop, t_period, seed, dt, ratio=None):
        """fix langevin Tstart Tstop damp seed [scale ...] (src/fix_langevin.cpp); call after set_atoms —
        the RanMars draws follow the order of that call — and before setup"""
        r = _d(ratio) if ratio is not None else None
        _chk(self.L.b200md_fix_langevin(self.h, C.c_double(t_start), C.c_double(t_stop), C.c_double(t_period),
                                        int(seed), _p(r), C.c_double(dt)))

    def fix_langevin_window(self, ntimestep, beginstep, endstep):
        _chk(self.L.b200md_fix_langevin_window(self.h, C.c_longlong(ntimestep), C.c_longlong(beginstep), C.c_longlong(endstep)))

    def fix_langevin_sort(self, sortfreq):
        """Atom::sortfreq of the host whose order the stream follows (default 1000, 0 = never re-sorted)"""
        _chk(self.L.b200md_fix_langevin_sort(self.h, int(sortfreq)))

    def fix_langevin_off(self):
        _chk(self.L.b200md_fix_langevin_off(self.h))

    def ranmars_stream(self, seed, n, fills=1):
        """`fills` consecutive device fills of n RanMars draws (the per-evaluation request of fix langevin)"""
        out = np.empty(n * fills, np.float64)
        _chk(self.L.b200md_ranmars_device_stream(self.h, int(seed), int(n), int(fills), _p(out)))
        return out

    def stress_atom(self, ke=True, which=PERATOM_PAIR | PERATOM_BOND):
        """`compute stress/atom NULL [virial]` (src/compute_stress_atom.cpp:255-340): -(m v v + W) * nktv2p, tag order"""
        _, w = self.peratom(0, 1, which)
        a = self.get_atoms(("tag", "v", "type"))
        if ke:
            m = self.mass[a["type"]][:, None] * self.units["mvv2e"]
            v = a["v"]
            w = w + m * np.stack([v[:, 0] * v[:, 0], v[:, 1] * v[:, 1], v[:, 2] * v[:, 2],
                                  v[:, 0] * v[:, 1], v[:, 0] * v[:, 2], v[:, 1] * v[:, 2]], axis=1)
        out = np.empty_like(w)
        out[a["tag"] - 1] = -self.units["nktv2p"] * w
        return out

    # fix nve / thermo / loop
    def nve_initial(self, dtv, dtf):
        _chk(self.L.b200md_nve_initial(self.h, C.c_double(dtv), C.c_double(dtf)))

    def nve_final(self, dtf):
        _chk(self.L.b200md_nve_final(self.h, C.c_double(dtf)))

    def thermo(self):
        t = Thermo()
        _chk(self.L.b200md_thermo_compute(self.h, C.byref(t)))
        return t.as_dict()

    def setup(self):
        _chk(self.L.b200md_setup(self.h))

    def run(self, nsteps, dt, thermo_every=0):
        cap = (nsteps // thermo_every + 2) if thermo_every > 0 else 2
        arr = (Thermo * cap)()
        n = C.c_int()
        _chk(self.L.b200md_run(self.h, int(nsteps), C.c_double(dt), int(thermo_every), arr, cap, C.byref(n)))
        return [arr[k].as_dict() for k in range(n.value)]

    def timing(self):
        t = Timing()
        _chk(self.L.b200md_get_timing(self.h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in Timing._fields_}

    def stream(self):
        s = C.c_void_p()
        _chk(self.L.b200md_get_stream(self.h, C.byref(s)))
        return s.value


# ---- PairLJCut host side (src/pair_lj_cut.cpp:442-556, src/pair.cpp:705-733) -------------------
def lj_cut_tables(ntypes, coeffs, cut_global, mix="geometric", shift=False):
    """coeffs: {(i,j): (epsilon, sigma[, cut])} for i<=j as given by pair_coeff; missing i<j pairs
    are mixed like Pair::mix_energy / mix_distance.  Returns 1-based (ntypes+1)^2 tables."""
    n1 = ntypes + 1
    eps = np.zeros((n1, n1)); sig = np.zeros((n1, n1)); cut = np.zeros((n1, n1)); setf = np.zeros((n1, n1), bool)
    for (i, j), c in coeffs.items():
        eps[i, j], sig[i, j] = c[0], c[1]
        cut[i, j] = c[2] if len(c) > 2 else cut_global
        setf[i, j] = True
    T = {k: np.zeros((n1, n1)) for k in ("cutsq", "lj1", "lj2", "lj3", "lj4", "offset")}
    for i in range(1, n1):
        for j in range(i, n1):
            if not setf[i, j]:
                if not (setf[i, i] and setf[j, j]):
                    raise B200MDError("All pair coeffs are not set")
                if mix == "geometric":
                    eps[i, j] = math.sqrt(eps[i, i] * eps[j, j]); sig[i, j] = math.sqrt(sig[i, i] * sig[j, j])
                elif mix == "arithmetic":
                    eps[i, j] = math.sqrt(eps[i, i] * eps[j, j]); sig[i, j] = 0.5 * (sig[i, i] + sig[j, j])
                else:
                    raise B200MDError("unsupported pair_modify mix " + mix)
                # PairLJCut::init_one: cut[i][j] = mix_distance(cut[i][i], cut[j][j]) (src/pair_lj_cut.cpp:506,
                # Pair::mix_distance src/pair.cpp:722-733) -- the same rule as sigma
                cut[i, j] = math.sqrt(cut[i, i] * cut[j, j]) if mix == "geometric" else 0.5 * (cut[i, i] + cut[j, j])
            e, s, c = eps[i, j], sig[i, j], cut[i, j]
            v = dict(cutsq=c * c, lj1=48.0 * e * s ** 12, lj2=24.0 * e * s ** 6, lj3=4.0 * e * s ** 12,
                     lj4=4.0 * e * s ** 6, offset=0.0)
            if shift and c > 0.0:
                ratio = s / c
                v["offset"] = 4.0 * e * (ratio ** 12 - ratio ** 6)
            for k in T:
                T[k][i, j] = T[k][j, i] = v[k]
    return T


def cutneighsq_table(ntypes, cutsq, skin):
    """Neighbor::init (src/neighbor.cpp:348-388): cutneighsq = (sqrt(cutsq)+skin)^2."""
    n1 = ntypes + 1
    out = np.zeros((n1, n1)); cmax = 0.0
    for i in range(1, n1):
        for j in range(1, n1):
            c = math.sqrt(cutsq[i, j]) + skin
            out[i, j] = c * c
            cmax = max(cmax, c)
    return out, cmax
