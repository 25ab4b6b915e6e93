"""The bench inputs of the reference restated as host-side set-up calls on a b200md Context:
`bench/in.lj` and `bench/in.eam` (/root/reference/bench/in.lj:1-31, in.eam:1-32), scaled by
-var x/y/z like the originals.  Synthetic fcc lattice data, no files except the EAM potential.
"""
import math
import os

import numpy as np

from . import b200md as B  # noqa: F401  (package import)


DATA_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data")


def data_file(name):
    """input data the bench configs name (Cu_u3.eam, Cu_mishin1.eam.alloy, data.chain): data/<name>, put there by
    __graft_entry__.build() from the reference tree (bench/, potentials/)"""
    p = os.path.join(DATA_DIR, name)
    if not os.path.exists(p):
        raise B.B200MDError("input data file %s is missing; run __graft_entry__.build() where the reference tree exists" % p)
    return p


def _decompose(ctx, boxlo, boxhi):
    prd = [boxhi[d] - boxlo[d] for d in range(3)]
    grid = B.procgrid(ctx.nranks, prd)
    ctx.set_box(boxlo, boxhi, (1, 1, 1), grid)
    return grid


def in_lj(ctx, x=1, y=1, z=1, list_kind=B.LIST_FULL, every=20, delay=0, check=0, temperature=1.44, seed=87287):
    """bench/in.lj: units lj; lattice fcc 0.8442; region box block 0 20x 0 20y 0 20z;
    create_atoms 1 box; mass 1 1.0; velocity all create 1.44 87287 loop geom;
    pair_style lj/cut 2.5; pair_coeff 1 1 1.0 1.0 2.5; neighbor 0.3 bin;
    neigh_modify delay 0 every 20 check no; fix 1 all nve."""
    ctx.set_units("lj")
    scale = math.pow(4.0 / 0.8442, 1.0 / 3.0)  # Lattice: (nbasis/volume/rho*)^(1/dim), src/lattice.cpp:256-262
    nx, ny, nz = 20 * x, 20 * y, 20 * z
    boxlo = [0.0, 0.0, 0.0]
    boxhi = [nx * scale, ny * scale, nz * scale]
    _decompose(ctx, boxlo, boxhi)
    sublo, subhi = ctx.subbox()
    mass = np.array([0.0, 1.0])
    xs, vs, tag = B.lattice_fcc(scale, nx, ny, nz, sublo, subhi, seed, 1.0)
    ctx.set_atoms(xs, vs, np.ones(len(tag), np.int32), tag, mass)
    ctx.velocity_finalize(temperature, True)
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 2.5)}, 2.5)
    ctx.pair_lj_cut_config(T["cutsq"], T["lj1"], T["lj2"], T["lj3"], T["lj4"], T["offset"])
    skin = 0.3
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], skin)
    ctx.neigh_config(cn, cmax, skin, every=every, delay=delay, check=check, list_kind=list_kind)
    return dict(natoms=4 * nx * ny * nz, dt=0.005, thermo=0, units="lj",
                neigh=dict(cutneighsq=cn, cutneighmax=cmax, skin=skin, every=every, delay=delay, check=check))


def in_eam(ctx, x=1, y=1, z=1, potential=None, style="eam", temperature=1600.0, seed=376847,
           list_kind=B.LIST_FULL):
    """bench/in.eam: units metal; lattice fcc 3.615; region box block 0 20x 0 20y 0 20z;
    create_atoms 1 box; pair_style eam; pair_coeff 1 1 Cu_u3.eam; velocity all create 1600.0
    376847 loop geom; neighbor 1.0 bin; neigh_modify every 1 delay 5 check yes; fix 1 all nve;
    timestep 0.005; thermo 50.  style="eam/alloy" takes a setfl file (Cu_mishin1.eam.alloy)."""
    from . import eam_host
    ctx.set_units("metal")
    scale = 3.615
    nx, ny, nz = 20 * x, 20 * y, 20 * z
    boxlo = [0.0, 0.0, 0.0]
    boxhi = [nx * scale, ny * scale, nz * scale]
    _decompose(ctx, boxlo, boxhi)
    sublo, subhi = ctx.subbox()
    if potential is None:
        potential = data_file("Cu_u3.eam" if style == "eam" else "Cu_mishin1.eam.alloy")
    if style == "eam":
        tb = eam_host.tables_from_funcfl([potential])
    else:
        tb = eam_host.tables_from_setfl(potential, ["Cu"])
    mass = np.array([0.0, tb["mass"][1]])
    xs, vs, tag = B.lattice_fcc(scale, nx, ny, nz, sublo, subhi, seed, mass[1])
    ctx.set_atoms(xs, vs, np.ones(len(tag), np.int32), tag, mass)
    ctx.velocity_finalize(temperature, True)
    ctx.pair_eam_config(tb)
    skin = 1.0
    cutsq = np.zeros((2, 2)); cutsq[1, 1] = tb["cutforcesq"]
    cn, cmax = B.cutneighsq_table(1, cutsq, skin)
    ctx.neigh_config(cn, cmax, skin, every=1, delay=5, check=1, list_kind=list_kind)
    return dict(natoms=4 * nx * ny * nz, dt=0.005, thermo=50, units="metal",
                neigh=dict(cutneighsq=cn, cutneighmax=cmax, skin=skin, every=1, delay=5, check=1))


def atom_sort_order(x, sublo, subhi, cutneighmax):
    """the local order Verlet::setup leaves (Atom::sort, src/atom.cpp:2246-2349, src/verlet.cpp:123)"""
    return B.atom_sort_order(x, sublo, subhi, cutneighmax)


def in_chain(ctx, data_path=None, x=1, y=1, z=1, list_kind=B.LIST_HALF_NEWTON, langevin=False):
    """bench/in.chain / in.chain.scaled (`replicate x y z`)
    (/root/reference/bench/in.chain:1-28, in.chain.scaled:1-33): units lj;
    atom_style bond; special_bonds fene; read_data data.chain; neighbor 0.4 bin;
    neigh_modify every 1 delay 1; bond_style fene; bond_coeff 1 30.0 1.5 1.0 1.0;
    pair_style lj/cut 1.12; pair_modify shift yes; pair_coeff 1 1 1.0 1.0 1.12; fix 1 all nve;
    timestep 0.012; and with langevin=True `fix 2 all langevin 1.0 1.0 10.0 904297`.
    fix langevin draws its RanMars numbers in local atom order (src/fix_langevin.cpp:636-662), so
    with langevin=True the atoms are uploaded in the order Atom::sort gives them at the reference's
    set-up (one rank only)."""
    from . import molecule_host as M
    if data_path is None:
        data_path = data_file("data.chain")
    d = M.replicate(M.read_data(data_path), x, y, z)
    ctx.set_units("lj")
    boxlo, boxhi = list(d["boxlo"]), list(d["boxhi"])
    _decompose(ctx, boxlo, boxhi)
    sublo, subhi = ctx.subbox()
    xs = d["x"]
    own = np.ones(len(xs), bool)
    for k in range(3):
        own &= (xs[:, k] >= sublo[k]) & (xs[:, k] < subhi[k])
    T = B.lj_cut_tables(1, {(1, 1): (1.0, 1.0, 1.12)}, 1.12, shift=True)
    skin = 0.4
    cn, cmax = B.cutneighsq_table(1, T["cutsq"], skin)
    sel = np.flatnonzero(own)
    if langevin:
        sel = d["local_order"][own[d["local_order"]]]        # the reference's local order before its set-up
        sel = sel[atom_sort_order(xs[sel], sublo, subhi, cmax)]
    ctx.set_atoms(xs[sel], d["v"][sel], d["type"][sel], d["tag"][sel], d["mass"], image=d["image"][sel])
    if langevin:
        ctx.fix_langevin(1.0, 1.0, 10.0, 904297, 0.012)
    tp = M.build_topology(d["natoms"], d["bonds"])          # special_bonds fene: lj = coul = 0 1 1
    ctx.set_topology(tp)
    ctx.bond_fene_config([0.0, 30.0], [0.0, 1.5], [0.0, 1.0], [0.0, 1.0])
    ctx.pair_lj_cut_config(T["cutsq"], T["lj1"], T["lj2"], T["lj3"], T["lj4"], T["offset"])
    ctx.neigh_config(cn, cmax, skin, every=1, delay=1, check=1, list_kind=list_kind)
    return dict(natoms=d["natoms"], dt=0.012, thermo=100, units="lj", data=d, topology=tp, tables=T,
                neigh=dict(cutneighsq=cn, cutneighmax=cmax, skin=skin, every=1, delay=1, check=1))
