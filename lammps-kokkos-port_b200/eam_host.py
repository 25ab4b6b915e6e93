"""Host side of pair_style eam / eam/alloy: potential-file readers and spline construction.

Mirrors (same names, same arithmetic order) the reference's one-time set-up:
  PairEAM::read_file        /root/reference/src/MANYBODY/pair_eam.cpp:460-526   (funcfl)
  PairEAM::file2array       pair_eam.cpp:533-740   (4-point Lagrange re-gridding, z2r=27.2*0.529*zi*zj)
  PairEAM::array2spline     pair_eam.cpp:744-765
  PairEAM::interpolate      pair_eam.cpp:769-797
  PairEAMAlloy::read_file   pair_eam_alloy.cpp:109-219 (setfl)
  PairEAMAlloy::file2array  pair_eam_alloy.cpp:225-337
  PairEAMFS::read_file      pair_eam_fs.cpp:111-220    (setfl with per-pair densities)
  PairEAMFS::file2array     pair_eam_fs.cpp:226-345
  unit conversion metal<->real (utils.cpp:1500-1510 factor 23.060549): applied like the readers do,
  loop bounds included (eam/alloy leaves the last table entry unconverted, pair_eam_alloy.cpp:161,169)
The product of this module is the table set handed to b200md_pair_eam_config.
"""
import numpy as np


def _tokens(path, skip):
    with open(path) as fh:
        lines = fh.read().split("\n")
    return lines[:skip], " ".join(lines[skip:]).split()


METAL2REAL = 23.060549


def _conv(units_file, units_run):
    if units_file == units_run:
        return None
    if units_file == "metal" and units_run == "real":
        return METAL2REAL
    if units_file == "real" and units_run == "metal":
        return 1.0 / METAL2REAL
    raise ValueError("unsupported unit conversion %s -> %s" % (units_file, units_run))


def read_funcfl(path, conv=None):
    with open(path) as fh:
        lines = fh.read().split("\n")
    l2 = lines[1].split()
    f = dict(mass=float(l2[1]))
    l3 = lines[2].split()
    f["nrho"] = int(l3[0]); f["drho"] = float(l3[1]); f["nr"] = int(l3[2])
    f["dr"] = float(l3[3]); f["cut"] = float(l3[4])
    if f["nrho"] <= 0 or f["nr"] <= 0 or f["dr"] <= 0.0:
        raise ValueError("Invalid EAM potential file")
    rest = np.array([float(t) for t in " ".join(lines[3:]).split()])
    nrho, nr = f["nrho"], f["nr"]
    # 1-based arrays like the reference (index 0 unused)
    f["frho"] = np.concatenate([[0.0], rest[:nrho]])
    f["zr"] = np.concatenate([[0.0], rest[nrho:nrho + nr]])
    f["rhor"] = np.concatenate([[0.0], rest[nrho + nr:nrho + 2 * nr]])
    if conv is not None:
        f["frho"][1:] *= conv
        f["zr"][1:] *= np.sqrt(conv)
    return f


def read_fs(path, conv=None):
    """DYNAMO setfl, Finnis-Sinclair form: rhor[i][j] per element pair."""
    _, tok = _tokens(path, 3)
    it = iter(tok)
    ne = int(next(it))
    s = dict(nelements=ne, elements=[next(it) for _ in range(ne)])
    s["nrho"] = int(next(it)); s["drho"] = float(next(it)); s["nr"] = int(next(it))
    s["dr"] = float(next(it)); s["cut"] = float(next(it))
    nrho, nr = s["nrho"], s["nr"]
    s["mass"] = []; s["frho"] = np.zeros((ne, nrho + 1)); s["rhor"] = np.zeros((ne, ne, nr + 1))
    s["z2r"] = np.zeros((ne, ne, nr + 1))
    for i in range(ne):
        next(it)
        s["mass"].append(float(next(it)))
        next(it); next(it)
        s["frho"][i, 1:] = [float(next(it)) for _ in range(nrho)]
        if conv is not None:
            s["frho"][i, 1:] *= conv
        for j in range(ne):
            s["rhor"][i, j, 1:] = [float(next(it)) for _ in range(nr)]
    for i in range(ne):
        for j in range(i + 1):
            s["z2r"][i, j, 1:] = [float(next(it)) for _ in range(nr)]
            if conv is not None:
                s["z2r"][i, j, 1:] *= conv
    return s


def read_setfl(path, conv=None):
    _, tok = _tokens(path, 3)
    it = iter(tok)
    ne = int(next(it))
    s = dict(nelements=ne, elements=[next(it) for _ in range(ne)])
    s["nrho"] = int(next(it)); s["drho"] = float(next(it)); s["nr"] = int(next(it))
    s["dr"] = float(next(it)); s["cut"] = float(next(it))
    nrho, nr = s["nrho"], s["nr"]
    s["mass"] = []; s["frho"] = np.zeros((ne, nrho + 1)); s["rhor"] = np.zeros((ne, nr + 1))
    s["z2r"] = np.zeros((ne, ne, nr + 1))
    for i in range(ne):
        next(it)
        s["mass"].append(float(next(it)))
        next(it); next(it)  # lattice constant, lattice type
        s["frho"][i, 1:] = [float(next(it)) for _ in range(nrho)]
        s["rhor"][i, 1:] = [float(next(it)) for _ in range(nr)]
        if conv is not None:
            s["frho"][i, 1:nrho] *= conv          # j = 1 .. nrho-1, as the reference does
    for i in range(ne):
        for j in range(i + 1):
            s["z2r"][i, j, 1:] = [float(next(it)) for _ in range(nr)]
            if conv is not None:
                s["z2r"][i, j, 1:nr] *= conv      # k = 1 .. nr-1
    return s


def _regrid(src, n_src, d_src, n_dst, d_dst):
    """4-point Lagrange interpolation of a 1-based table onto (m-1)*d_dst, m=1..n_dst
    (pair_eam.cpp:577-597)."""
    sixth = 1.0 / 6.0
    m = np.arange(1, n_dst + 1)
    r = (m - 1) * d_dst
    p = r / d_src + 1.0
    k = p.astype(np.int64)
    k = np.minimum(k, n_src - 2)
    k = np.maximum(k, 2)
    p = p - k
    p = np.minimum(p, 2.0)
    cof1 = -sixth * p * (p - 1.0) * (p - 2.0)
    cof2 = 0.5 * (p * p - 1.0) * (p - 2.0)
    cof3 = -0.5 * p * (p + 1.0) * (p - 2.0)
    cof4 = sixth * p * (p * p - 1.0)
    out = np.zeros(n_dst + 1)
    out[1:] = cof1 * src[k - 1] + cof2 * src[k] + cof3 * src[k + 1] + cof4 * src[k + 2]
    return out


def interpolate(n, delta, f):
    """PairEAM::interpolate: 1-based [n+1][7] spline block from the 1-based table f."""
    sp = np.zeros((n + 1, 7))
    sp[1:n + 1, 6] = f[1:n + 1]
    sp[1, 5] = sp[2, 6] - sp[1, 6]
    sp[2, 5] = 0.5 * (sp[3, 6] - sp[1, 6])
    sp[n - 1, 5] = 0.5 * (sp[n, 6] - sp[n - 2, 6])
    sp[n, 5] = sp[n, 6] - sp[n - 1, 6]
    m = np.arange(3, n - 1)
    sp[m, 5] = ((sp[m - 2, 6] - sp[m + 2, 6]) + 8.0 * (sp[m + 1, 6] - sp[m - 1, 6])) / 12.0
    m = np.arange(1, n)
    sp[m, 4] = 3.0 * (sp[m + 1, 6] - sp[m, 6]) - 2.0 * sp[m, 5] - sp[m + 1, 5]
    sp[m, 3] = sp[m, 5] + sp[m + 1, 5] - 2.0 * (sp[m + 1, 6] - sp[m, 6])
    sp[n, 4] = 0.0
    sp[n, 3] = 0.0
    m = np.arange(1, n + 1)
    sp[m, 2] = sp[m, 5] / delta
    sp[m, 1] = 2.0 * sp[m, 4] / delta
    sp[m, 0] = 3.0 * sp[m, 3] / delta
    return sp


def _type_maps(ntypes, map_, fs_nelements=0):
    n1 = ntypes + 1
    type2rhor = np.zeros((n1, n1), np.int32); type2z2r = np.zeros((n1, n1), np.int32)
    for i in range(1, n1):
        for j in range(1, n1):
            # eam, eam/alloy: density of the SOURCE atom's element (pair_eam.cpp:672,
            # pair_eam_alloy.cpp:296); eam/fs: per ordered element pair (pair_eam_fs.cpp:307)
            type2rhor[i, j] = map_[i] * fs_nelements + map_[j] if fs_nelements else map_[i]
            irow, icol = map_[i], map_[j]
            if irow == -1 or icol == -1:
                type2z2r[i, j] = 0
                continue
            if irow < icol:
                irow, icol = map_[j], map_[i]
            n = sum(mm + 1 for mm in range(irow)) + icol
            type2z2r[i, j] = n
    return type2rhor, type2z2r


def _finish(ntypes, map_, nrho, nr, drho, dr, rhomax, cutmax, frho, rhor, z2r, masses, fs_nelements=0):
    nfrho = frho.shape[0]
    type2frho = np.zeros(ntypes + 1, np.int32)
    for i in range(1, ntypes + 1):
        type2frho[i] = map_[i] if map_[i] >= 0 else nfrho - 1
    type2rhor, type2z2r = _type_maps(ntypes, map_, fs_nelements)
    tb = dict(nrho=nrho, nr=nr, rdr=1.0 / dr, rdrho=1.0 / drho, rhomax=rhomax, cutforcesq=cutmax * cutmax,
              cutmax=cutmax,
              frho_spline=np.stack([interpolate(nrho, drho, frho[i]) for i in range(nfrho)]),
              rhor_spline=np.stack([interpolate(nr, dr, rhor[i]) for i in range(rhor.shape[0])]),
              z2r_spline=np.stack([interpolate(nr, dr, z2r[i]) for i in range(z2r.shape[0])]),
              type2frho=type2frho, type2rhor=type2rhor, type2z2r=type2z2r,
              scale=np.ones((ntypes + 1, ntypes + 1)), mass=np.array([0.0] + masses))
    return tb


def tables_from_funcfl(files_by_type, units="metal"):
    """pair_style eam; pair_coeff i i file_i  (files_by_type[i-1] is the funcfl file of type i)."""
    conv = _conv("metal", units)
    uniq = []
    map_ = [-1]
    for p in files_by_type:
        if p not in uniq:
            uniq.append(p)
        map_.append(uniq.index(p))
    ntypes = len(files_by_type)
    F = [read_funcfl(p, conv) for p in uniq]
    dr = max(f["dr"] for f in F); drho = max(f["drho"] for f in F)
    rmax = max((f["nr"] - 1) * f["dr"] for f in F); rhomax = max((f["nrho"] - 1) * f["drho"] for f in F)
    nr = int(rmax / dr + 0.5); nrho = int(rhomax / drho + 0.5)
    nf = len(F)
    frho = np.zeros((nf + 1, nrho + 1))
    for i, f in enumerate(F):
        frho[i] = _regrid(f["frho"], f["nrho"], f["drho"], nrho, drho)
    rhor = np.stack([_regrid(f["rhor"], f["nr"], f["dr"], nr, dr) for f in F])
    z2r = []
    for i in range(nf):
        for j in range(i + 1):
            zri = _regrid(F[i]["zr"], F[i]["nr"], F[i]["dr"], nr, dr)
            zrj = _regrid(F[j]["zr"], F[j]["nr"], F[j]["dr"], nr, dr)
            z2r.append(27.2 * 0.529 * zri * zrj)
    cutmax = max(f["cut"] for f in F)
    masses = [F[map_[i]]["mass"] for i in range(1, ntypes + 1)]
    return _finish(ntypes, map_, nrho, nr, drho, dr, rhomax, cutmax, frho, rhor, np.stack(z2r), masses)


def _element_map(s, elements_by_type):
    map_ = [-1]
    for e in elements_by_type:
        if e is None:
            map_.append(-1)
        else:
            if e not in s["elements"]:
                raise ValueError("No matching element in EAM potential file")
            map_.append(s["elements"].index(e))
    return map_


def tables_from_fs(path, elements_by_type, units="metal"):
    """pair_style eam/fs; pair_coeff * * file E1 E2 ..."""
    s = read_fs(path, _conv("metal", units))
    ntypes = len(elements_by_type)
    map_ = _element_map(s, elements_by_type)
    ne, nrho, nr = s["nelements"], s["nrho"], s["nr"]
    frho = np.zeros((ne + 1, nrho + 1)); frho[:ne] = s["frho"]
    rhor = s["rhor"].reshape(ne * ne, nr + 1)
    z2r = np.stack([s["z2r"][i, j] for i in range(ne) for j in range(i + 1)])
    masses = [s["mass"][map_[i]] if map_[i] >= 0 else 1.0 for i in range(1, ntypes + 1)]
    return _finish(ntypes, map_, nrho, nr, s["drho"], s["dr"], (nrho - 1) * s["drho"], s["cut"], frho, rhor,
                   z2r, masses, fs_nelements=ne)


def tables_from_setfl(path, elements_by_type, units="metal"):
    """pair_style eam/alloy; pair_coeff * * file E1 E2 ...  (NULL -> None)."""
    s = read_setfl(path, _conv("metal", units))
    ntypes = len(elements_by_type)
    map_ = _element_map(s, elements_by_type)
    ne, nrho, nr = s["nelements"], s["nrho"], s["nr"]
    frho = np.zeros((ne + 1, nrho + 1)); frho[:ne] = s["frho"]
    z2r = np.stack([s["z2r"][i, j] for i in range(ne) for j in range(i + 1)])
    masses = [s["mass"][map_[i]] if map_[i] >= 0 else 1.0 for i in range(1, ntypes + 1)]
    return _finish(ntypes, map_, nrho, nr, s["drho"], s["dr"], (nrho - 1) * s["drho"], s["cut"], frho, s["rhor"],
                   z2r, masses)
