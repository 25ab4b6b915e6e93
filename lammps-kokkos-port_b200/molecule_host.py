"""Host side of `atom_style bond` inputs (bench/in.chain): data-file reader, special-neighbor
construction and the tag-keyed CSR tables handed to b200md_set_topology.

Mirrors the reference's one-time set-up:
  ReadData (atoms / velocities / bonds sections)   /root/reference/src/read_data.cpp
  Atom::data_bonds    src/atom.cpp:1408-1470   newton_bond on: a bond is stored on its first atom
  Special::build      src/special.cpp:54-120   1-2 from the bonds (both directions); 1-3 / 1-4 only
                                               when their weights differ from 1.0 (early return)
  Special::dedup / combine                     an atom listed at a closer level is dropped from
                                               the farther ones; self is never listed
  special_bonds fene  src/input.cpp (special_bonds): lj = coul = (0, 1, 1)
"""
import numpy as np


def read_data(path):
    """LAMMPS data file with Atoms (atom_style bond: id mol type x y z [ix iy iz]), Velocities,
    Bonds, Masses.  Returns a dict of numpy arrays ordered by tag."""
    with open(path) as fh:
        lines = fh.read().split("\n")
    hdr = {}
    box = {}
    sections = {}
    i = 1
    names = ("Masses", "Atoms", "Velocities", "Bonds", "Bond Coeffs", "Pair Coeffs")
    while i < len(lines):
        ln = lines[i].split("#")[0].strip()
        if not ln:
            i += 1
            continue
        if ln in names:
            j = i + 2
            rows = []
            while j < len(lines) and lines[j].strip():
                rows.append(lines[j].split())
                j += 1
            sections[ln] = rows
            i = j
            continue
        tok = ln.split()
        if len(tok) >= 2 and tok[-1] in ("atoms", "bonds", "angles", "dihedrals", "impropers"):
            hdr[tok[-1]] = int(tok[0])
        elif len(tok) == 3 and tok[-1] == "types":
            hdr[tok[1] + " types"] = int(tok[0])
        elif len(tok) == 4 and tok[2].endswith("lo"):
            box[tok[2][0]] = (float(tok[0]), float(tok[1]))
        i += 1
    n = hdr["atoms"]
    x = np.zeros((n, 3)); v = np.zeros((n, 3)); typ = np.zeros(n, np.int32); mol = np.zeros(n, np.int32)
    img = np.zeros((n, 3), np.int32)
    for r in sections["Atoms"]:
        k = int(r[0]) - 1
        mol[k] = int(r[1]); typ[k] = int(r[2]); x[k] = [float(t) for t in r[3:6]]
        if len(r) >= 9:
            img[k] = [int(t) for t in r[6:9]]
    for r in sections.get("Velocities", []):
        v[int(r[0]) - 1] = [float(t) for t in r[1:4]]
    nt = hdr["atom types"]
    mass = np.zeros(nt + 1)
    for r in sections["Masses"]:
        mass[int(r[0])] = float(r[1])
    bonds = np.array([[int(r[1]), int(r[2]), int(r[3])] for r in sections.get("Bonds", [])], dtype=np.int64).reshape(-1, 3)
    IMG = 512
    image = ((img[:, 0] + IMG) & 1023) | (((img[:, 1] + IMG) & 1023) << 10) | (((img[:, 2] + IMG) & 1023) << 20)
    return dict(natoms=n, boxlo=np.array([box[k][0] for k in "xyz"]), boxhi=np.array([box[k][1] for k in "xyz"]),
                x=x, v=v, type=typ, molecule=mol, image=image.astype(np.int32), mass=mass,
                tag=np.arange(1, n + 1, dtype=np.int32), bonds=bonds, nbondtypes=hdr.get("bond types", 0))


def replicate(d, nx, ny, nz):
    """`replicate nx ny nz` (src/replicate.cpp:146-150, 430-640): unmap by the image flags, copy
    into every replica cell (loops ix, iy, iz), remap into the enlarged box, tags / bond partners /
    molecule ids offset by (iz*ny*nx + iy*nx + ix) * maxtag.  Arrays come back sorted by tag;
    `local_order` indexes them in the reference's local (creation) order — replica after replica,
    the old atoms in their order inside each — which is what order-dependent host semantics
    (Atom::sort's stable bins, hence fix langevin's draws) start from."""
    if (nx, ny, nz) == (1, 1, 1):
        out = dict(d)
        out["local_order"] = np.arange(d["natoms"])
        return out
    n = d["natoms"]
    prd = d["boxhi"] - d["boxlo"]
    img = d["image"]
    ib = np.stack([(img & 1023) - 512, ((img >> 10) & 1023) - 512, ((img >> 20) & 1023) - 512], axis=1)
    xu = d["x"] + ib * prd                                      # Domain::unmap
    newhi = d["boxlo"] + prd * np.array([nx, ny, nz])
    newprd = newhi - d["boxlo"]
    maxtag = int(d["tag"].max()); maxmol = int(d["molecule"].max())
    X, V, T, TAG, MOL, IMG, BONDS = [], [], [], [], [], [], []
    for ix in range(nx):
        for iy in range(ny):
            for iz in range(nz):
                off = iz * ny * nx + iy * nx + ix
                x = xu + prd * np.array([ix, iy, iz])
                k = np.floor((x - d["boxlo"]) / newprd).astype(np.int64)   # Domain::remap
                x = x - k * newprd
                # guard the half-open interval against rounding at the upper face
                hi = x >= newhi
                x[hi] -= np.broadcast_to(newprd, x.shape)[hi]; k[hi] += 1
                lo = x < d["boxlo"]
                x[lo] += np.broadcast_to(newprd, x.shape)[lo]; k[lo] -= 1
                X.append(x); V.append(d["v"]); T.append(d["type"]); TAG.append(d["tag"] + off * maxtag)
                MOL.append(d["molecule"] + off * maxmol)
                IMG.append(((k[:, 0] + 512) & 1023) | (((k[:, 1] + 512) & 1023) << 10) | (((k[:, 2] + 512) & 1023) << 20))
                b = d["bonds"].copy(); b[:, 1:] += off * maxtag
                BONDS.append(b)
    tag = np.concatenate(TAG)
    o = np.argsort(tag, kind="stable")
    out = dict(d)
    out.update(natoms=n * nx * ny * nz, boxhi=newhi, x=np.concatenate(X)[o], v=np.concatenate(V)[o],
               type=np.concatenate(T)[o].astype(np.int32), tag=tag[o].astype(np.int32),
               molecule=np.concatenate(MOL)[o].astype(np.int32), image=np.concatenate(IMG)[o].astype(np.int32),
               bonds=np.concatenate(BONDS), local_order=np.argsort(o, kind="stable"))
    return out


def _csr(natoms, owner, cols):
    """rows keyed by 1-based tag; offsets have natoms+2 entries; stable in input order"""
    order = np.argsort(owner, kind="stable")
    counts = np.bincount(owner, minlength=natoms + 1)
    off = np.zeros(natoms + 2, np.int64)
    off[1:] = np.cumsum(counts)
    return off.astype(np.int32), [np.ascontiguousarray(c[order]).astype(np.int32) for c in cols]


def build_topology(natoms, bonds, special_lj=(0.0, 0.0, 1.0, 1.0), special_coul=(0.0, 0.0, 1.0, 1.0)):
    """bonds: [type, atom1, atom2] rows.  Returns the CSR tables of b200md_set_topology."""
    btype, a1, a2 = bonds[:, 0], bonds[:, 1], bonds[:, 2]
    bond_off, (bond_partner, bond_type) = _csr(natoms, a1, [a2, btype])
    # 1-2: both directions
    own = np.concatenate([a1, a2]); oth = np.concatenate([a2, a1])
    need_far = not (special_lj[2] == 1.0 and special_coul[2] == 1.0 and special_lj[3] == 1.0 and special_coul[3] == 1.0)
    if not need_far:                                      # Special::build returns after the 1-2 pass
        spec_off, (spec_tag, spec_lev) = _csr(natoms, own, [oth, np.ones(len(own), np.int64)])
        return dict(natoms=natoms, bond_offset=bond_off, bond_partner=bond_partner, bond_type=bond_type,
                    special_offset=spec_off, special_tag=spec_tag, special_level=spec_lev,
                    special_lj=np.array(special_lj, dtype=np.float64), special_coul=np.array(special_coul, dtype=np.float64))
    sets = [dict() for _ in range(natoms + 1)]            # tag -> {partner tag: level}
    for o, p in zip(own.tolist(), oth.tolist()):
        sets[o].setdefault(p, 1)
    if need_far:
        one2 = [list(s.keys()) for s in sets]
        for level in (2, 3):
            prev = [[p for p, l in s.items() if l == level - 1] for s in sets]
            for t in range(1, natoms + 1):
                for p in prev[t]:
                    for q in one2[p]:
                        if q != t:
                            sets[t].setdefault(q, level)
    own, tags, lev = [], [], []
    for t in range(1, natoms + 1):
        for p, l in sets[t].items():
            own.append(t); tags.append(p); lev.append(l)
    own = np.array(own, dtype=np.int64); tags = np.array(tags, dtype=np.int64); lev = np.array(lev, dtype=np.int64)
    # levels ascending within a row, like Atom::special (1-2 first)
    o = np.lexsort((lev, own))
    spec_off, (spec_tag, spec_lev) = _csr(natoms, own[o], [tags[o], lev[o]])
    return dict(natoms=natoms, bond_offset=bond_off, bond_partner=bond_partner, bond_type=bond_type,
                special_offset=spec_off, special_tag=spec_tag, special_level=spec_lev,
                special_lj=np.array(special_lj, dtype=np.float64), special_coul=np.array(special_coul, dtype=np.float64))
