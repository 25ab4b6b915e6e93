/* md_oracle.c — TEST INFRASTRUCTURE ONLY.  A plain-C, single-process CPU restatement of the
 * reference's short-range hot path, used by tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg as the CHECKER.  The product (libb200md.so) never links or calls this.
 *
 * Parity status: PINNED — tests/test_oracle_vs_reference.py checks every function below
 * against the unmodified reference compiled into oracle/_ref/liblammps_ref.so
 * (oracle/build_ref.py) and against the thermo printed in the reference's own logs
 * (bench/log.6Oct16.{lj,eam}.fixed.icc.1), and tests/golden/ holds vectors generated from
 * that reference build (tests/golden/make_golden.py).
 *
 * Each function cites the reference code (paths relative to /root/reference/src) it follows:
 * same operand order, no FMA contraction (compile with -ffp-contract=off), same comparison
 * operators, linked-list bins, 6-swap periodic borders, half/newton and full lists.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define MAXSWAP 30
#define BIG 1.0e20
#define SMALL 1.0e-6
#define MIN(a, b) ((a) < (b) ? (a) : (b))
#define MAX(a, b) ((a) > (b) ? (a) : (b))
#define IMGMASK 1023
#define IMGBITS 10

typedef struct {
  /* units (update.cpp:145-200) */
  double boltz, mvv2e, ftm2v, nktv2p;
  /* domain */
  double boxlo[3], boxhi[3], prd[3];
  /* atoms: owned 0..nlocal-1, ghosts nlocal..nlocal+nghost-1 */
  int nlocal, nghost, nmax, ntypes;
  double *x, *v, *f; /* [nmax][3] */
  int *type, *tag, *image;
  int *owner; /* ghost -> owned index it images (for rho/f reverse, fp forward on one proc) */
  double *mass;
  /* neighbor */
  double *cutneighsq; /* (ntypes+1)^2 */
  double cutneighmax, skin, triggersq;
  int every, delay, check, half;
  int ago, nbuilds, ndanger;
  double *xhold;
  int nbinx, nbiny, nbinz, mbinx, mbiny, mbinz, mbinxlo, mbinylo, mbinzlo, mbins;
  double binsizex, binsizey, binsizez, bininvx, bininvy, bininvz;
  int *binhead, *bins, *atom2bin;
  int nstencil, *stencil;
  int *numneigh, *firstoff, *neigh;
  long neighcap;
  /* swaps (6, one proc, periodic) */
  int nswap, sendnum[MAXSWAP], firstrecv[MAXSWAP], *sendlist[MAXSWAP], sendcap[MAXSWAP], pbcdim[MAXSWAP], pbcsign[MAXSWAP];
  /* pair */
  int pair_style; /* 1 lj, 2 eam */
  double *cutsq, *lj1, *lj2, *lj3, *lj4, *offset;
  int nrho, nr, nfrho, nrhor, nz2r;
  double rdr, rdrho, rhomax, cutforcesq;
  double *frho_spline, *rhor_spline, *z2r_spline, *scale;
  int *type2frho, *type2rhor, *type2z2r;
  double *rho, *fp;
  double eng_vdwl, virial[6];
  /* molecular topology keyed by tag (CSR, rows 1..natoms) + bond_style fene */
  int molecular, topo_natoms;
  int *bond_off, *bond_partner, *bond_type, *spec_off, *spec_tag, *spec_level;
  int special_flag[4];
  int *map, *sametag;          /* Atom::map_set (atom_map.cpp:220-300): tag -> lowest local index, image chain */
  int nbondlist, maxbond, *bondlist; /* [n][3] i1,i2,type */
  int nbondtypes;
  double *fene_k, *fene_r0, *fene_eps, *fene_sigma;
  double eng_bond, virial_bond[6];
  int bond_error;
  /* per-atom tallies (Pair::ev_tally / Bond::ev_tally eflag_atom, vflag_atom branches) */
  int eflag_atom, vflag_atom;
  double *eatom, *vatom; /* [nmax], [nmax][6] pair + bond, as compute pe/atom / stress/atom sum them */
  long step;
  /* Atom::sort schedule (atom.cpp:102-103, 2252): 0 = the caller never sorted, sorting stays off */
  long sortfreq, nextsort;
  int in_run;
  /* fix langevin (fix_langevin.cpp) with its RanMars stream (random_mars.cpp) */
  int langevin;
  double lang_tstart, lang_tstop, lang_period, lang_dt, *lang_ratio;
  long lang_begin, lang_end;
  double mars_u[98], mars_c, mars_cd, mars_cm;
  int mars_i97, mars_j97;
} ora_sim;

typedef struct {
  long step;
  double natoms, temp, pe, ke, press, virial[6], emol;
} ora_thermo;

static void grow(ora_sim *s, int need) {
  if (need <= s->nmax) return;
  int n = need + need / 4 + 1024;
  s->x = realloc(s->x, sizeof(double) * 3 * n);
  s->v = realloc(s->v, sizeof(double) * 3 * n);
  s->f = realloc(s->f, sizeof(double) * 3 * n);
  s->xhold = realloc(s->xhold, sizeof(double) * 3 * n);
  s->type = realloc(s->type, sizeof(int) * n);
  s->tag = realloc(s->tag, sizeof(int) * n);
  s->image = realloc(s->image, sizeof(int) * n);
  s->owner = realloc(s->owner, sizeof(int) * n);
  s->bins = realloc(s->bins, sizeof(int) * n);
  s->atom2bin = realloc(s->atom2bin, sizeof(int) * n);
  s->numneigh = realloc(s->numneigh, sizeof(int) * n);
  s->firstoff = realloc(s->firstoff, sizeof(int) * n);
  s->rho = realloc(s->rho, sizeof(double) * n);
  s->fp = realloc(s->fp, sizeof(double) * n);
  s->eatom = realloc(s->eatom, sizeof(double) * n);
  s->vatom = realloc(s->vatom, sizeof(double) * 6 * n);
  s->nmax = n;
}

ora_sim *ora_create(const double *boxlo, const double *boxhi, int n, const double *x, const double *v,
                    const int *type, const int *tag, int ntypes, const double *mass, double boltz,
                    double mvv2e, double ftm2v, double nktv2p) {
  ora_sim *s = calloc(1, sizeof(ora_sim));
  s->boltz = boltz; s->mvv2e = mvv2e; s->ftm2v = ftm2v; s->nktv2p = nktv2p;
  for (int d = 0; d < 3; d++) { s->boxlo[d] = boxlo[d]; s->boxhi[d] = boxhi[d]; s->prd[d] = boxhi[d] - boxlo[d]; }
  s->ntypes = ntypes;
  s->mass = malloc(sizeof(double) * (ntypes + 1));
  memcpy(s->mass, mass, sizeof(double) * (ntypes + 1));
  grow(s, n * 2);
  s->nlocal = n;
  memcpy(s->x, x, sizeof(double) * 3 * n);
  if (v) memcpy(s->v, v, sizeof(double) * 3 * n); else memset(s->v, 0, sizeof(double) * 3 * n);
  memset(s->f, 0, sizeof(double) * 3 * n);
  memcpy(s->type, type, sizeof(int) * n);
  memcpy(s->tag, tag, sizeof(int) * n);
  for (int i = 0; i < n; i++) s->image[i] = (512 << 20) | (512 << 10) | 512;
  return s;
}

void ora_destroy(ora_sim *s) {
  if (!s) return;
  free(s->x); free(s->v); free(s->f); free(s->xhold); free(s->type); free(s->tag); free(s->image); free(s->owner);
  free(s->bins); free(s->atom2bin); free(s->numneigh); free(s->firstoff); free(s->rho); free(s->fp); free(s->eatom); free(s->vatom); free(s->mass);
  free(s->cutneighsq); free(s->binhead); free(s->stencil); free(s->neigh);
  for (int k = 0; k < MAXSWAP; k++) free(s->sendlist[k]);
  free(s->cutsq); free(s->lj1); free(s->lj2); free(s->lj3); free(s->lj4); free(s->offset);
  free(s->frho_spline); free(s->rhor_spline); free(s->z2r_spline); free(s->scale);
  free(s->type2frho); free(s->type2rhor); free(s->type2z2r); free(s->lang_ratio);
  free(s);
}

/* Velocity::zero_momentum + rescale with ComputeTemp (velocity.cpp:358-368; compute_temp.cpp:72-99) */
static double temp_scalar(ora_sim *s) {
  double t = 0.0;
  for (int i = 0; i < s->nlocal; i++) {
    double *v = &s->v[3 * i];
    t += (v[0] * v[0] + v[1] * v[1] + v[2] * v[2]) * s->mass[s->type[i]];
  }
  double dof = 3.0 * s->nlocal - 3.0;
  double tfactor = dof > 0 ? s->mvv2e / (dof * s->boltz) : 0.0;
  return t * tfactor;
}
void ora_velocity_finalize(ora_sim *s, double t_desired) {
  double p[3] = {0, 0, 0}, masstotal = 0;
  for (int i = 0; i < s->nlocal; i++) {
    double m = s->mass[s->type[i]];
    for (int d = 0; d < 3; d++) p[d] += s->v[3 * i + d] * m;
    masstotal += m;
  }
  for (int d = 0; d < 3; d++) p[d] /= masstotal;
  for (int i = 0; i < s->nlocal; i++)
    for (int d = 0; d < 3; d++) s->v[3 * i + d] -= p[d];
  double t = temp_scalar(s);
  double factor = sqrt(t_desired / t);
  for (int i = 0; i < 3 * s->nlocal; i++) s->v[i] *= factor;
}

/* Domain::pbc (domain.cpp:530-650), orthogonal, fully periodic */
static void pbc(ora_sim *s) {
  for (int i = 0; i < s->nlocal; i++)
    for (int d = 0; d < 3; d++) {
      double *xc = &s->x[3 * i + d];
      int sh = d * IMGBITS;
      if (*xc < s->boxlo[d]) {
        *xc += s->prd[d];
        int idim = (s->image[i] >> sh) & IMGMASK, other = s->image[i] ^ (idim << sh);
        idim--; idim &= IMGMASK;
        s->image[i] = other | (idim << sh);
      }
      if (*xc >= s->boxhi[d]) {
        *xc -= s->prd[d];
        *xc = MAX(*xc, s->boxlo[d]);
        int idim = (s->image[i] >> sh) & IMGMASK, other = s->image[i] ^ (idim << sh);
        idim++; idim &= IMGMASK;
        s->image[i] = other | (idim << sh);
      }
    }
}

/* CommBrick::setup + borders on one periodic process (comm_brick.cpp:412-482, 771-979;
 * AtomVec::pack_border atom_vec.cpp:794-830): 2*maxneed self-swaps per dimension,
 * maxneed = int(cutghost/prd)+1; swaps after the first pair of a dimension forward only the
 * ghosts that arrived in the previous pair and cut their slab at the box middle. */
static int borders(ora_sim *s) {
  const double cut = s->cutneighmax;
  s->nghost = 0;
  int iswap = 0;
  for (int dim = 0; dim < 3; dim++) {
    const int maxneed = (int)(cut * 1 / s->prd[dim]) + 1;
    if (iswap + 2 * maxneed > MAXSWAP) return -1;
    const double mid = 0.5 * (s->boxlo[dim] + s->boxhi[dim]);
    int nlast = 0, nfirst = 0;
    for (int ineed = 0; ineed < 2 * maxneed; ineed++, iswap++) {
      double lo, hi;
      int sign;
      if (ineed % 2 == 0) {
        lo = ineed < 2 ? -BIG : mid; hi = s->boxlo[dim] + cut; sign = 1;
        nfirst = nlast; nlast = s->nlocal + s->nghost;
      } else { lo = s->boxhi[dim] - cut; hi = ineed < 2 ? BIG : mid; sign = -1; }
      int nsend = 0;
      for (int i = nfirst; i < nlast; i++)
        if (s->x[3 * i + dim] >= lo && s->x[3 * i + dim] <= hi) {
          if (nsend == s->sendcap[iswap]) {
            s->sendcap[iswap] = s->sendcap[iswap] * 2 + 1024;
            s->sendlist[iswap] = realloc(s->sendlist[iswap], sizeof(int) * s->sendcap[iswap]);
          }
          s->sendlist[iswap][nsend++] = i;
        }
      int first = s->nlocal + s->nghost;
      grow(s, first + nsend);
      for (int k = 0; k < nsend; k++) {
        int j = s->sendlist[iswap][k], g = first + k;
        for (int d = 0; d < 3; d++) s->x[3 * g + d] = s->x[3 * j + d];
        s->x[3 * g + dim] = s->x[3 * j + dim] + sign * s->prd[dim];
        s->tag[g] = s->tag[j]; s->type[g] = s->type[j];
        s->owner[g] = j < s->nlocal ? j : s->owner[j];
      }
      s->sendnum[iswap] = nsend; s->firstrecv[iswap] = first;
      s->pbcdim[iswap] = dim; s->pbcsign[iswap] = sign;
      s->nghost += nsend;
    }
  }
  s->nswap = iswap;
  return 0;
}
/* CommBrick::forward_comm (comm_brick.cpp:538-591) + AtomVec::pack_comm (atom_vec.cpp:349-378) */
static void forward_comm(ora_sim *s) {
  for (int iswap = 0; iswap < s->nswap; iswap++)
    for (int k = 0; k < s->sendnum[iswap]; k++) {
      int j = s->sendlist[iswap][k], g = s->firstrecv[iswap] + k, dim = s->pbcdim[iswap];
      for (int d = 0; d < 3; d++) s->x[3 * g + d] = s->x[3 * j + d];
      s->x[3 * g + dim] = s->x[3 * j + dim] + s->pbcsign[iswap] * s->prd[dim];
    }
}
/* CommBrick::reverse_comm (comm_brick.cpp:598-639): swaps in reverse order, += into senders */
static void reverse_comm_vec(ora_sim *s, double *a, int nv) {
  for (int iswap = s->nswap - 1; iswap >= 0; iswap--)
    for (int k = 0; k < s->sendnum[iswap]; k++) {
      int j = s->sendlist[iswap][k], g = s->firstrecv[iswap] + k;
      for (int d = 0; d < nv; d++) a[nv * j + d] += a[nv * g + d];
    }
}
static void forward_comm_scalar(ora_sim *s, double *a) {
  for (int iswap = 0; iswap < s->nswap; iswap++)
    for (int k = 0; k < s->sendnum[iswap]; k++) a[s->firstrecv[iswap] + k] = a[s->sendlist[iswap][k]];
}

/* NBinStandard::setup_bins (nbin_standard.cpp:82-214), one process: sub-box = box */
static void setup_bins(ora_sim *s) {
  double bbox[3], lo[3], hi[3];
  for (int d = 0; d < 3; d++) { bbox[d] = s->prd[d]; lo[d] = s->boxlo[d] - s->cutneighmax; hi[d] = s->boxhi[d] + s->cutneighmax; }
  double binsize_optimal = 0.5 * s->cutneighmax;
  double binsizeinv = 1.0 / binsize_optimal;
  int nb[3]; double bs[3], bi[3]; int mlo[3], mhi[3];
  for (int d = 0; d < 3; d++) {
    nb[d] = (int)(bbox[d] * binsizeinv);
    if (nb[d] == 0) nb[d] = 1;
    bs[d] = bbox[d] / nb[d];
    bi[d] = 1.0 / bs[d];
    double coord = lo[d] - SMALL * bbox[d];
    mlo[d] = (int)((coord - s->boxlo[d]) * bi[d]);
    if (coord < s->boxlo[d]) mlo[d] = mlo[d] - 1;
    coord = hi[d] + SMALL * bbox[d];
    mhi[d] = (int)((coord - s->boxlo[d]) * bi[d]);
    mlo[d] -= 1; mhi[d] += 1;
  }
  s->nbinx = nb[0]; s->nbiny = nb[1]; s->nbinz = nb[2];
  s->binsizex = bs[0]; s->binsizey = bs[1]; s->binsizez = bs[2];
  s->bininvx = bi[0]; s->bininvy = bi[1]; s->bininvz = bi[2];
  s->mbinxlo = mlo[0]; s->mbinylo = mlo[1]; s->mbinzlo = mlo[2];
  s->mbinx = mhi[0] - mlo[0] + 1; s->mbiny = mhi[1] - mlo[1] + 1; s->mbinz = mhi[2] - mlo[2] + 1;
  s->mbins = s->mbinx * s->mbiny * s->mbinz;
  s->binhead = realloc(s->binhead, sizeof(int) * (s->mbins + 1));
}

/* NBin::coord2bin (nbin.cpp:139-168) */
static int coord2bin(const ora_sim *s, const double *x) {
  int ix, iy, iz;
  if (x[0] >= s->boxhi[0]) ix = (int)((x[0] - s->boxhi[0]) * s->bininvx) + s->nbinx;
  else if (x[0] >= s->boxlo[0]) { ix = (int)((x[0] - s->boxlo[0]) * s->bininvx); ix = MIN(ix, s->nbinx - 1); }
  else ix = (int)((x[0] - s->boxlo[0]) * s->bininvx) - 1;
  if (x[1] >= s->boxhi[1]) iy = (int)((x[1] - s->boxhi[1]) * s->bininvy) + s->nbiny;
  else if (x[1] >= s->boxlo[1]) { iy = (int)((x[1] - s->boxlo[1]) * s->bininvy); iy = MIN(iy, s->nbiny - 1); }
  else iy = (int)((x[1] - s->boxlo[1]) * s->bininvy) - 1;
  if (x[2] >= s->boxhi[2]) iz = (int)((x[2] - s->boxhi[2]) * s->bininvz) + s->nbinz;
  else if (x[2] >= s->boxlo[2]) { iz = (int)((x[2] - s->boxlo[2]) * s->bininvz); iz = MIN(iz, s->nbinz - 1); }
  else iz = (int)((x[2] - s->boxlo[2]) * s->bininvz) - 1;
  return (iz - s->mbinzlo) * s->mbiny * s->mbinx + (iy - s->mbinylo) * s->mbinx + (ix - s->mbinxlo);
}

/* NBinStandard::bin_atoms (nbin_standard.cpp:221-260): linked lists, ghosts end up last */
static void bin_atoms(ora_sim *s) {
  int nall = s->nlocal + s->nghost;
  for (int i = 0; i < s->mbins; i++) s->binhead[i] = -1;
  for (int i = nall - 1; i >= 0; i--) {
    int ibin = coord2bin(s, &s->x[3 * i]);
    s->atom2bin[i] = ibin;
    s->bins[i] = s->binhead[ibin];
    s->binhead[ibin] = i;
  }
}

/* NStencil::bin_distance (nstencil.cpp:405-422) */
static double bin_distance(const ora_sim *s, int i, int j, int k) {
  double delx, dely, delz;
  if (i > 0) delx = (i - 1) * s->binsizex; else if (i == 0) delx = 0.0; else delx = (i + 1) * s->binsizex;
  if (j > 0) dely = (j - 1) * s->binsizey; else if (j == 0) dely = 0.0; else dely = (j + 1) * s->binsizey;
  if (k > 0) delz = (k - 1) * s->binsizez; else if (k == 0) delz = 0.0; else delz = (k + 1) * s->binsizez;
  return delx * delx + dely * dely + delz * delz;
}
/* NStencil::create_setup (nstencil.cpp:213-247) + NStencilFullBin3d::create
 * (nstencil_full_bin_3d.cpp:26-37) / NStencilHalfBin3d::create (nstencil_half_bin_3d.cpp:26-38) */
static void create_stencil(ora_sim *s) {
  int sx = (int)(s->cutneighmax * s->bininvx); if (sx * s->binsizex < s->cutneighmax) sx++;
  int sy = (int)(s->cutneighmax * s->bininvy); if (sy * s->binsizey < s->cutneighmax) sy++;
  int sz = (int)(s->cutneighmax * s->bininvz); if (sz * s->binsizez < s->cutneighmax) sz++;
  double cutsq = s->cutneighmax * s->cutneighmax;
  s->stencil = realloc(s->stencil, sizeof(int) * (2 * sx + 1) * (2 * sy + 1) * (2 * sz + 1));
  s->nstencil = 0;
  if (!s->half) {
    for (int k = -sz; k <= sz; k++)
      for (int j = -sy; j <= sy; j++)
        for (int i = -sx; i <= sx; i++)
          if (bin_distance(s, i, j, k) < cutsq) s->stencil[s->nstencil++] = k * s->mbiny * s->mbinx + j * s->mbinx + i;
  } else {
    for (int k = 0; k <= sz; k++)
      for (int j = -sy; j <= sy; j++)
        for (int i = -sx; i <= sx; i++)
          if (k > 0 || j > 0 || (j == 0 && i > 0))
            if (bin_distance(s, i, j, k) < cutsq) s->stencil[s->nstencil++] = k * s->mbiny * s->mbinx + j * s->mbinx + i;
  }
}

static void push_neigh(ora_sim *s, long *n, int j) {
  if (*n == s->neighcap) { s->neighcap = s->neighcap * 2 + (1 << 20); s->neigh = realloc(s->neigh, sizeof(int) * s->neighcap); }
  s->neigh[(*n)++] = j;
}

/* NPair::find_special (npair.h:125-160) on the tag-keyed special table: 0 keep, -1 exclude
 * (weights with flag 2 are not restated: ora_set_topology refuses them) */
static int find_special(const ora_sim *s, int itag, int jtag) {
  for (int k = s->spec_off[itag]; k < s->spec_off[itag + 1]; k++)
    if (s->spec_tag[k] == jtag) return s->special_flag[s->spec_level[k]] == 0 ? -1 : 0;
  return 0;
}
/* Domain::minimum_image_check (domain.h:165-171), fully periodic box */
static int minimum_image_check(const ora_sim *s, double dx, double dy, double dz) {
  if (fabs(dx) > 0.5 * s->prd[0]) return 1;
  if (fabs(dy) > 0.5 * s->prd[1]) return 1;
  if (fabs(dz) > 0.5 * s->prd[2]) return 1;
  return 0;
}
/* the molecular branch of NPairHalfBinNewton::build (npair_half_bin_newton.cpp:98-113) */
static void push_pair(ora_sim *s, long *n, int i, int j, double delx, double dely, double delz) {
  if (s->molecular) {
    int which = find_special(s, s->tag[i], s->tag[j]);
    if (which == 0) push_neigh(s, n, j);
    else if (minimum_image_check(s, delx, dely, delz)) push_neigh(s, n, j);
    return;
  }
  push_neigh(s, n, j);
}

/* NPairFullBinAtomonly::build (npair_full_bin_atomonly.cpp:32-92),
 * NPairHalfBinAtomonlyNewton::build (npair_half_bin_atomonly_newton.cpp:33-113) and, with a
 * topology, NPairHalfBinNewton::build (npair_half_bin_newton.cpp:36-161) */
static void build_list(ora_sim *s) {
  const int nlocal = s->nlocal, nt1 = s->ntypes + 1;
  const double *x = s->x;
  long n = 0;
  for (int i = 0; i < nlocal; i++) {
    long start = n;
    int itype = s->type[i];
    double xtmp = x[3 * i], ytmp = x[3 * i + 1], ztmp = x[3 * i + 2];
    if (s->half) {
      for (int j = s->bins[i]; j >= 0; j = s->bins[j]) {
        if (j >= nlocal) {
          if (x[3 * j + 2] < ztmp) continue;
          if (x[3 * j + 2] == ztmp) {
            if (x[3 * j + 1] < ytmp) continue;
            if (x[3 * j + 1] == ytmp && x[3 * j] < xtmp) continue;
          }
        }
        double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
        double rsq = delx * delx + dely * dely + delz * delz;
        if (rsq <= s->cutneighsq[itype * nt1 + s->type[j]]) push_pair(s, &n, i, j, delx, dely, delz);
      }
    }
    int ibin = s->atom2bin[i];
    for (int k = 0; k < s->nstencil; k++)
      for (int j = s->binhead[ibin + s->stencil[k]]; j >= 0; j = s->bins[j]) {
        if (!s->half && i == j) continue;
        double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
        double rsq = delx * delx + dely * dely + delz * delz;
        if (rsq <= s->cutneighsq[itype * nt1 + s->type[j]]) push_pair(s, &n, i, j, delx, dely, delz);
      }
    s->firstoff[i] = (int)start;
    s->numneigh[i] = (int)(n - start);
  }
}

void ora_neigh_config(ora_sim *s, const double *cutneighsq, double cutneighmax, double skin, int every, int delay,
                      int check, int half) {
  int nt = (s->ntypes + 1) * (s->ntypes + 1);
  s->cutneighsq = realloc(s->cutneighsq, sizeof(double) * nt);
  memcpy(s->cutneighsq, cutneighsq, sizeof(double) * nt);
  s->cutneighmax = cutneighmax; s->skin = skin; s->triggersq = 0.25 * skin * skin; /* neighbor.cpp:348 */
  s->every = every; s->delay = delay; s->check = check; s->half = half;
  setup_bins(s);
  create_stencil(s);
}

static void map_set(ora_sim *s);
static int bond_list_build(ora_sim *s);
static void atom_sort(ora_sim *s);
/* the reneighbor branch of Verlet::run (verlet.cpp:289-307) + Neighbor::build (neighbor.cpp:2398-2492) */
int ora_neigh_build(ora_sim *s) {
  pbc(s);
  if (s->in_run && s->sortfreq > 0 && s->step >= s->nextsort) atom_sort(s); /* verlet.cpp:285 */
  if (borders(s)) return -1;
  for (int i = 0; i < 3 * s->nlocal; i++) s->xhold[i] = s->x[i];
  bin_atoms(s);
  build_list(s);
  if (s->molecular) { map_set(s); if (bond_list_build(s)) return -1; }
  s->ago = 0;
  s->nbuilds++;
  return 0;
}

/* Neighbor::decide + check_distance (neighbor.cpp:2309-2390) */
static int decide(ora_sim *s) {
  s->ago++;
  if (s->ago >= s->delay && s->ago % s->every == 0) {
    if (!s->check) return 1;
    int flag = 0;
    for (int i = 0; i < s->nlocal; i++) {
      double delx = s->x[3 * i] - s->xhold[3 * i], dely = s->x[3 * i + 1] - s->xhold[3 * i + 1],
             delz = s->x[3 * i + 2] - s->xhold[3 * i + 2];
      double rsq = delx * delx + dely * dely + delz * delz;
      if (rsq > s->triggersq) flag = 1;
    }
    if (flag && s->ago == MAX(s->every, s->delay)) s->ndanger++;
    return flag;
  }
  return 0;
}

/* ---- pair styles ------------------------------------------------------------------------ */
static double *dup_d(const double *p, long n) { double *q = malloc(sizeof(double) * n); memcpy(q, p, sizeof(double) * n); return q; }
static int *dup_i(const int *p, long n) { int *q = malloc(sizeof(int) * n); memcpy(q, p, sizeof(int) * n); return q; }

void ora_pair_lj(ora_sim *s, const double *cutsq, const double *lj1, const double *lj2, const double *lj3,
                 const double *lj4, const double *offset) {
  long nt = (s->ntypes + 1) * (s->ntypes + 1);
  s->cutsq = dup_d(cutsq, nt); s->lj1 = dup_d(lj1, nt); s->lj2 = dup_d(lj2, nt); s->lj3 = dup_d(lj3, nt);
  s->lj4 = dup_d(lj4, nt); s->offset = dup_d(offset, nt);
  s->pair_style = 1;
}
void ora_pair_eam(ora_sim *s, int nrho, int nr, double rdr, double rdrho, double rhomax, double cutforcesq, int nfrho,
                  const double *frho_spline, int nrhor, const double *rhor_spline, int nz2r, const double *z2r_spline,
                  const int *type2frho, const int *type2rhor, const int *type2z2r, const double *scale) {
  long nt = (s->ntypes + 1) * (s->ntypes + 1);
  s->nrho = nrho; s->nr = nr; s->rdr = rdr; s->rdrho = rdrho; s->rhomax = rhomax; s->cutforcesq = cutforcesq;
  s->nfrho = nfrho; s->nrhor = nrhor; s->nz2r = nz2r;
  s->frho_spline = dup_d(frho_spline, (long)nfrho * (nrho + 1) * 7);
  s->rhor_spline = dup_d(rhor_spline, (long)nrhor * (nr + 1) * 7);
  s->z2r_spline = dup_d(z2r_spline, (long)nz2r * (nr + 1) * 7);
  s->type2frho = dup_i(type2frho, s->ntypes + 1); s->type2rhor = dup_i(type2rhor, nt); s->type2z2r = dup_i(type2z2r, nt);
  s->scale = dup_d(scale, nt);
  s->pair_style = 2;
}

/* Pair::virial_fdotr_compute (pair.cpp:1735-1782): sum over owned + ghost before reverse_comm */
static void virial_fdotr(ora_sim *s) {
  int nall = s->nlocal + s->nghost;
  for (int k = 0; k < 6; k++) s->virial[k] = 0.0;
  for (int i = 0; i < nall; i++) {
    const double *f = &s->f[3 * i], *x = &s->x[3 * i];
    s->virial[0] += f[0] * x[0]; s->virial[1] += f[1] * x[1]; s->virial[2] += f[2] * x[2];
    s->virial[3] += f[1] * x[0]; s->virial[4] += f[2] * x[0]; s->virial[5] += f[2] * x[1];
  }
}

/* eflag_atom / vflag_atom branches of Pair::ev_tally (pair.cpp:1036-1040, 1078-1097) and of
 * Bond::ev_tally (bond.cpp:186-192, 234-251) with newton on: half of the pair's energy and of
 * v = del (x) del * fpair to each of the two atoms, ghosts included (reverse-communicated later) */
static void tally_atom(ora_sim *s, int i, int j, double e, double fpair, double delx, double dely, double delz) {
  if (s->eflag_atom) { s->eatom[i] += 0.5 * e; s->eatom[j] += 0.5 * e; }
  if (s->vflag_atom) {
    double v[6] = {delx * delx * fpair, dely * dely * fpair, delz * delz * fpair,
                   delx * dely * fpair, delx * delz * fpair, dely * delz * fpair};
    for (int k = 0; k < 6; k++) { s->vatom[6 * i + k] += 0.5 * v[k]; s->vatom[6 * j + k] += 0.5 * v[k]; }
  }
}

/* PairLJCut::compute (pair_lj_cut.cpp:69-137) on the half list with newton_pair on */
static void compute_lj(ora_sim *s, int eflag, int vflag) {
  const int nt1 = s->ntypes + 1;
  double *x = s->x, *f = s->f;
  s->eng_vdwl = 0.0;
  for (int i = 0; i < s->nlocal; i++) {
    double xtmp = x[3 * i], ytmp = x[3 * i + 1], ztmp = x[3 * i + 2];
    int itype = s->type[i];
    const int *jlist = &s->neigh[s->firstoff[i]];
    for (int jj = 0; jj < s->numneigh[i]; jj++) {
      int j = jlist[jj];
      double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
      double rsq = delx * delx + dely * dely + delz * delz;
      int ij = itype * nt1 + s->type[j];
      if (rsq < s->cutsq[ij]) {
        double r2inv = 1.0 / rsq;
        double r6inv = r2inv * r2inv * r2inv;
        double forcelj = r6inv * (s->lj1[ij] * r6inv - s->lj2[ij]);
        double fpair = forcelj * r2inv;
        f[3 * i] += delx * fpair; f[3 * i + 1] += dely * fpair; f[3 * i + 2] += delz * fpair;
        f[3 * j] -= delx * fpair; f[3 * j + 1] -= dely * fpair; f[3 * j + 2] -= delz * fpair;
        double evdwl = r6inv * (s->lj3[ij] * r6inv - s->lj4[ij]) - s->offset[ij];
        if (eflag) s->eng_vdwl += evdwl;
        if (s->eflag_atom || s->vflag_atom) tally_atom(s, i, j, evdwl, fpair, delx, dely, delz);
      }
    }
  }
  if (vflag) virial_fdotr(s);
}

/* PairEAM::compute (MANYBODY/pair_eam.cpp:142-328), half list, newton_pair on */
static void compute_eam(ora_sim *s, int eflag, int vflag) {
  const int nt1 = s->ntypes + 1, nlocal = s->nlocal, nall = s->nlocal + s->nghost;
  double *x = s->x, *f = s->f, *rho = s->rho, *fp = s->fp;
  const int nr = s->nr, nrho = s->nrho;
  s->eng_vdwl = 0.0;
  for (int i = 0; i < nall; i++) rho[i] = 0.0;
  for (int i = 0; i < nlocal; i++) { /* :187-220 */
    double xtmp = x[3 * i], ytmp = x[3 * i + 1], ztmp = x[3 * i + 2];
    int itype = s->type[i];
    const int *jlist = &s->neigh[s->firstoff[i]];
    for (int jj = 0; jj < s->numneigh[i]; jj++) {
      int j = jlist[jj];
      double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
      double rsq = delx * delx + dely * dely + delz * delz;
      if (rsq < s->cutforcesq) {
        int jtype = s->type[j];
        double p = sqrt(rsq) * s->rdr + 1.0;
        int m = (int)p;
        m = MIN(m, nr - 1);
        p -= m;
        p = MIN(p, 1.0);
        const double *coeff = &s->rhor_spline[((long)s->type2rhor[jtype * nt1 + itype] * (nr + 1) + m) * 7];
        rho[i] += ((coeff[3] * p + coeff[4]) * p + coeff[5]) * p + coeff[6];
        coeff = &s->rhor_spline[((long)s->type2rhor[itype * nt1 + jtype] * (nr + 1) + m) * 7];
        rho[j] += ((coeff[3] * p + coeff[4]) * p + coeff[5]) * p + coeff[6];
      }
    }
  }
  reverse_comm_vec(s, rho, 1); /* :224 */
  for (int i = 0; i < nlocal; i++) { /* :231-247 */
    double p = rho[i] * s->rdrho + 1.0;
    int m = (int)p;
    m = MAX(1, MIN(m, nrho - 1));
    p -= m;
    p = MIN(p, 1.0);
    const double *coeff = &s->frho_spline[((long)s->type2frho[s->type[i]] * (nrho + 1) + m) * 7];
    fp[i] = (coeff[0] * p + coeff[1]) * p + coeff[2];
    if (eflag) {
      double phi = ((coeff[3] * p + coeff[4]) * p + coeff[5]) * p + coeff[6];
      if (rho[i] > s->rhomax) phi += fp[i] * (rho[i] - s->rhomax);
      phi *= s->scale[s->type[i] * nt1 + s->type[i]];
      s->eng_vdwl += phi;
      if (s->eflag_atom) s->eatom[i] += phi; /* pair_eam.cpp:245 */
    }
  }
  forward_comm_scalar(s, fp); /* :251 */
  for (int i = 0; i < nlocal; i++) { /* :257-325 */
    double xtmp = x[3 * i], ytmp = x[3 * i + 1], ztmp = x[3 * i + 2];
    int itype = s->type[i];
    const int *jlist = &s->neigh[s->firstoff[i]];
    for (int jj = 0; jj < s->numneigh[i]; jj++) {
      int j = jlist[jj];
      double delx = xtmp - x[3 * j], dely = ytmp - x[3 * j + 1], delz = ztmp - x[3 * j + 2];
      double rsq = delx * delx + dely * dely + delz * delz;
      if (rsq < s->cutforcesq) {
        int jtype = s->type[j];
        double r = sqrt(rsq);
        double p = r * s->rdr + 1.0;
        int m = (int)p;
        m = MIN(m, nr - 1);
        p -= m;
        p = MIN(p, 1.0);
        const double *coeff = &s->rhor_spline[((long)s->type2rhor[itype * nt1 + jtype] * (nr + 1) + m) * 7];
        double rhoip = (coeff[0] * p + coeff[1]) * p + coeff[2];
        coeff = &s->rhor_spline[((long)s->type2rhor[jtype * nt1 + itype] * (nr + 1) + m) * 7];
        double rhojp = (coeff[0] * p + coeff[1]) * p + coeff[2];
        coeff = &s->z2r_spline[((long)s->type2z2r[itype * nt1 + jtype] * (nr + 1) + m) * 7];
        double z2p = (coeff[0] * p + coeff[1]) * p + coeff[2];
        double z2 = ((coeff[3] * p + coeff[4]) * p + coeff[5]) * p + coeff[6];
        double recip = 1.0 / r;
        double phi = z2 * recip;
        double phip = z2p * recip - phi * recip;
        double psip = fp[i] * rhojp + fp[j] * rhoip + phip;
        double fpair = -s->scale[itype * nt1 + jtype] * psip * recip;
        f[3 * i] += delx * fpair; f[3 * i + 1] += dely * fpair; f[3 * i + 2] += delz * fpair;
        f[3 * j] -= delx * fpair; f[3 * j + 1] -= dely * fpair; f[3 * j + 2] -= delz * fpair;
        if (eflag) s->eng_vdwl += s->scale[itype * nt1 + jtype] * phi;
        if (s->eflag_atom || s->vflag_atom) tally_atom(s, i, j, s->scale[itype * nt1 + jtype] * phi, fpair, delx, dely, delz);
      }
    }
  }
  if (vflag) virial_fdotr(s);
}

/* ---- molecular topology, bond_style fene -------------------------------------------------- */
int ora_set_topology(ora_sim *s, int natoms, const int *bond_off, const int *bond_partner, const int *bond_type,
                     const int *spec_off, const int *spec_tag, const int *spec_level, const double *special_lj,
                     const double *special_coul) {
  s->topo_natoms = natoms;
  s->bond_off = dup_i(bond_off, natoms + 2); s->bond_partner = dup_i(bond_partner, bond_off[natoms + 1]);
  s->bond_type = dup_i(bond_type, bond_off[natoms + 1]);
  s->spec_off = dup_i(spec_off, natoms + 2); s->spec_tag = dup_i(spec_tag, spec_off[natoms + 1]);
  s->spec_level = dup_i(spec_level, spec_off[natoms + 1]);
  for (int k = 1; k <= 3; k++) { /* Neighbor::init special_flag (neighbor.cpp:430-452) */
    if (special_lj[k] == 0.0 && special_coul[k] == 0.0) s->special_flag[k] = 0;
    else if (special_lj[k] == 1.0 && special_coul[k] == 1.0) s->special_flag[k] = 1;
    else return -1;
  }
  s->map = malloc(sizeof(int) * (natoms + 1));
  s->molecular = 1;
  return 0;
}
void ora_bond_fene(ora_sim *s, int nbondtypes, const double *k, const double *r0, const double *eps, const double *sigma) {
  s->nbondtypes = nbondtypes;
  s->fene_k = dup_d(k, nbondtypes + 1); s->fene_r0 = dup_d(r0, nbondtypes + 1);
  s->fene_eps = dup_d(eps, nbondtypes + 1); s->fene_sigma = dup_d(sigma, nbondtypes + 1);
}
/* Atom::map_set, array style (atom_map.cpp:220-260): loop from the last atom down so the map
 * holds the lowest index (owned before ghost) and sametag chains to the next image */
static void map_set(ora_sim *s) {
  int nall = s->nlocal + s->nghost;
  s->sametag = realloc(s->sametag, sizeof(int) * (nall + 1));
  for (int t = 0; t <= s->topo_natoms; t++) s->map[t] = -1;
  for (int i = nall - 1; i >= 0; i--) { s->sametag[i] = s->map[s->tag[i]]; s->map[s->tag[i]] = i; }
}
/* Domain::closest_image (domain.cpp:1123-1151) */
static int closest_image(const ora_sim *s, int i, int j) {
  if (j < 0) return j;
  const double *xi = &s->x[3 * i];
  int closest = j;
  double delx = xi[0] - s->x[3 * j], dely = xi[1] - s->x[3 * j + 1], delz = xi[2] - s->x[3 * j + 2];
  double rsqmin = delx * delx + dely * dely + delz * delz;
  while (s->sametag[j] >= 0) {
    j = s->sametag[j];
    delx = xi[0] - s->x[3 * j]; dely = xi[1] - s->x[3 * j + 1]; delz = xi[2] - s->x[3 * j + 2];
    double rsq = delx * delx + dely * dely + delz * delz;
    if (rsq < rsqmin) { rsqmin = rsq; closest = j; }
  }
  return closest;
}
/* NTopoBondAll::build (ntopo_bond_all.cpp:37-80), newton_bond on */
static int bond_list_build(ora_sim *s) {
  s->nbondlist = 0;
  for (int i = 0; i < s->nlocal; i++) {
    int t = s->tag[i];
    for (int m = s->bond_off[t]; m < s->bond_off[t + 1]; m++) {
      int atom1 = s->map[s->bond_partner[m]];
      if (atom1 == -1) return -1; /* "Bond atoms missing" */
      atom1 = closest_image(s, i, atom1);
      if (s->nbondlist == s->maxbond) { s->maxbond = s->maxbond * 2 + 4096; s->bondlist = realloc(s->bondlist, sizeof(int) * 3 * s->maxbond); }
      int *b = &s->bondlist[3 * s->nbondlist++];
      b[0] = i; b[1] = atom1; b[2] = s->bond_type[m];
    }
  }
  return 0;
}
/* BondFENE::compute (MOLECULE/bond_fene.cpp:53-129) + Bond::ev_tally (bond.cpp:174-256), newton_bond on */
static void compute_fene(ora_sim *s, int eflag, int vflag) {
  const double cuberoot2 = 1.25992104989487316476;
  s->eng_bond = 0; for (int k = 0; k < 6; k++) s->virial_bond[k] = 0;
  for (int n = 0; n < s->nbondlist; n++) {
    int i1 = s->bondlist[3 * n], i2 = s->bondlist[3 * n + 1], type = s->bondlist[3 * n + 2];
    double delx = s->x[3 * i1] - s->x[3 * i2], dely = s->x[3 * i1 + 1] - s->x[3 * i2 + 1], delz = s->x[3 * i1 + 2] - s->x[3 * i2 + 2];
    double rsq = delx * delx + dely * dely + delz * delz;
    double r0sq = s->fene_r0[type] * s->fene_r0[type];
    double rlogarg = 1.0 - rsq / r0sq;
    if (rlogarg < 0.1) { if (rlogarg <= -3.0) s->bond_error = 1; rlogarg = 0.1; }
    double fbond = -s->fene_k[type] / rlogarg, sr6 = 0.0, ebond = 0.0;
    if (rsq < cuberoot2 * s->fene_sigma[type] * s->fene_sigma[type]) {
      double sr2 = s->fene_sigma[type] * s->fene_sigma[type] / rsq;
      sr6 = sr2 * sr2 * sr2;
      fbond += 48.0 * s->fene_eps[type] * sr6 * (sr6 - 0.5) / rsq;
    }
    if (eflag) {
      ebond = -0.5 * s->fene_k[type] * r0sq * log(rlogarg);
      if (rsq < cuberoot2 * s->fene_sigma[type] * s->fene_sigma[type])
        ebond += 4.0 * s->fene_eps[type] * sr6 * (sr6 - 1.0) + s->fene_eps[type];
    }
    s->f[3 * i1] += delx * fbond; s->f[3 * i1 + 1] += dely * fbond; s->f[3 * i1 + 2] += delz * fbond;
    s->f[3 * i2] -= delx * fbond; s->f[3 * i2 + 1] -= dely * fbond; s->f[3 * i2 + 2] -= delz * fbond;
    if (eflag) s->eng_bond += ebond;
    if (s->eflag_atom || s->vflag_atom) tally_atom(s, i1, i2, ebond, fbond, delx, dely, delz);
    if (vflag) {
      s->virial_bond[0] += delx * delx * fbond; s->virial_bond[1] += dely * dely * fbond; s->virial_bond[2] += delz * delz * fbond;
      s->virial_bond[3] += delx * dely * fbond; s->virial_bond[4] += delx * delz * fbond; s->virial_bond[5] += dely * delz * fbond;
    }
  }
}

/* Verlet::force_clear + pair->compute + comm->reverse_comm (verlet.cpp:311-330) */
int ora_compute(ora_sim *s, int eflag, int vflag) {
  if (!s->half) return -1; /* the reference's pair styles consume the half list */
  int nall = s->nlocal + s->nghost;
  memset(s->f, 0, sizeof(double) * 3 * nall);
  if (s->pair_style == 1) compute_lj(s, eflag, vflag);
  else if (s->pair_style == 2) compute_eam(s, eflag, vflag);
  else return -1;
  if (s->molecular && s->fene_k) compute_fene(s, eflag, vflag);
  reverse_comm_vec(s, s->f, 3);
  return s->bond_error ? -1 : 0;
}

/* what `compute pe/atom` and `compute stress/atom NULL virial` add up and reverse-communicate
 * (compute_pe_atom.cpp:92-170, compute_stress_atom.cpp:140-240): pair + bond per-atom energy and
 * virial of the current configuration.  vatom is the raw sum; stress/atom prints -nktv2p * it. */
int ora_compute_peratom(ora_sim *s, int eflag_atom, int vflag_atom) {
  int nall = s->nlocal + s->nghost;
  memset(s->eatom, 0, sizeof(double) * nall);
  memset(s->vatom, 0, sizeof(double) * 6 * nall);
  s->eflag_atom = eflag_atom; s->vflag_atom = vflag_atom;
  int r = ora_compute(s, 1, 1);
  s->eflag_atom = s->vflag_atom = 0;
  reverse_comm_vec(s, s->eatom, 1);
  reverse_comm_vec(s, s->vatom, 6);
  return r;
}
double *ora_eatom(ora_sim *s) { return s->eatom; }
double *ora_vatom(ora_sim *s) { return s->vatom; }

/* FixNVE::initial_integrate / final_integrate (fix_nve.cpp:64-141) */
static void nve_initial(ora_sim *s, double dtv, double dtf) {
  for (int i = 0; i < s->nlocal; i++) {
    double dtfm = dtf / s->mass[s->type[i]];
    for (int d = 0; d < 3; d++) {
      s->v[3 * i + d] += dtfm * s->f[3 * i + d];
      s->x[3 * i + d] += dtv * s->v[3 * i + d];
    }
  }
}
static void nve_final(ora_sim *s, double dtf) {
  for (int i = 0; i < s->nlocal; i++) {
    double dtfm = dtf / s->mass[s->type[i]];
    for (int d = 0; d < 3; d++) s->v[3 * i + d] += dtfm * s->f[3 * i + d];
  }
}

/* ComputeTemp / ComputePE / ComputePressure (compute_temp.cpp:72-99, compute_pe.cpp:84-112,
 * compute_pressure.cpp:228-262) */
void ora_thermo_compute(ora_sim *s, ora_thermo *t) {
  double dof = 3.0 * s->nlocal - 3.0;
  double temp = temp_scalar(s);
  double vol = s->prd[0] * s->prd[1] * s->prd[2];
  t->step = s->step; t->natoms = s->nlocal; t->temp = temp; t->pe = s->eng_vdwl + s->eng_bond; t->emol = s->eng_bond;
  t->ke = temp * 0.5 * dof * s->boltz;
  for (int k = 0; k < 6; k++) t->virial[k] = s->virial[k] + s->virial_bond[k];
  t->press = (dof * s->boltz * temp + t->virial[0] + t->virial[1] + t->virial[2]) / 3.0 * (1.0 / vol) * s->nktv2p;
}

/* ---- Atom::sort (atom.cpp:2246-2349) with Atom::setup_sort_bins (atom.cpp:2351-2410) --------
 * Spatial sort of the owned atoms, bins of 1/2 cutneighmax over the (sub-)box, stable inside a
 * bin.  Verlet::setup calls it (verlet.cpp:123, sortfreq = 1000 by default) after Domain::pbc and
 * before the borders, so it fixes the local index order — which is what the order of a half
 * list's same-bin pairs and the RanMars draws of fix langevin follow. */
static void atom_sort(ora_sim *s) {
  if (s->sortfreq <= 0) s->sortfreq = 1000;
  s->nextsort = (s->step / s->sortfreq) * s->sortfreq + s->sortfreq;
  s->nghost = 0;
  double binsize = 0.5 * s->cutneighmax;
  if (binsize == 0.0) return;
  double bininv = 1.0 / binsize;
  int nb[3];
  double binv[3];
  for (int d = 0; d < 3; d++) {
    nb[d] = (int)((s->boxhi[d] - s->boxlo[d]) * bininv);
    if (nb[d] == 0) nb[d] = 1;
    binv[d] = nb[d] / (s->boxhi[d] - s->boxlo[d]);
  }
  int nbins = nb[0] * nb[1] * nb[2], n = s->nlocal;
  if (nbins == 1) return;
  int *head = malloc(sizeof(int) * nbins), *next = malloc(sizeof(int) * (n + 1)), *permute = malloc(sizeof(int) * (n + 1));
  for (int i = 0; i < nbins; i++) head[i] = -1;
  for (int i = n - 1; i >= 0; i--) {
    int c[3];
    for (int d = 0; d < 3; d++) {
      c[d] = (int)((s->x[3 * i + d] - s->boxlo[d]) * binv[d]);
      c[d] = MAX(c[d], 0);
      c[d] = MIN(c[d], nb[d] - 1);
    }
    int ibin = c[2] * nb[1] * nb[0] + c[1] * nb[0] + c[0];
    next[i] = head[ibin];
    head[ibin] = i;
  }
  int m = 0;
  for (int b = 0; b < nbins; b++)
    for (int i = head[b]; i >= 0; i = next[i]) permute[m++] = i;
  /* new atom I = old atom permute[I] (the reference does it in place, cycle by cycle; same result) */
  double *x = malloc(sizeof(double) * 3 * n), *v = malloc(sizeof(double) * 3 * n), *f = malloc(sizeof(double) * 3 * n);
  int *type = malloc(sizeof(int) * n), *tag = malloc(sizeof(int) * n), *image = malloc(sizeof(int) * n);
  for (int i = 0; i < n; i++) {
    int j = permute[i];
    for (int d = 0; d < 3; d++) { x[3 * i + d] = s->x[3 * j + d]; v[3 * i + d] = s->v[3 * j + d]; f[3 * i + d] = s->f[3 * j + d]; }
    type[i] = s->type[j]; tag[i] = s->tag[j]; image[i] = s->image[j];
  }
  memcpy(s->x, x, sizeof(double) * 3 * n); memcpy(s->v, v, sizeof(double) * 3 * n); memcpy(s->f, f, sizeof(double) * 3 * n);
  memcpy(s->type, type, sizeof(int) * n); memcpy(s->tag, tag, sizeof(int) * n); memcpy(s->image, image, sizeof(int) * n);
  free(x); free(v); free(f); free(type); free(tag); free(image); free(head); free(next); free(permute);
}

/* Verlet::setup's sort (verlet.cpp:123); it also switches on the in-run re-sort of Verlet::run
 * (verlet.cpp:285: at the first reneighboring with ntimestep >= nextsort) */
void ora_atom_sort(ora_sim *s) {
  pbc(s);
  atom_sort(s);
}
void ora_set_sortfreq(ora_sim *s, long sortfreq) { s->sortfreq = sortfreq; }

/* ---- RanMars (random_mars.cpp:30-95): Marsaglia's lagged-Fibonacci generator --------------- */
static double mars_uniform(ora_sim *s) {
  double uni = s->mars_u[s->mars_i97] - s->mars_u[s->mars_j97];
  if (uni < 0.0) uni += 1.0;
  s->mars_u[s->mars_i97] = uni;
  s->mars_i97--;
  if (s->mars_i97 == 0) s->mars_i97 = 97;
  s->mars_j97--;
  if (s->mars_j97 == 0) s->mars_j97 = 97;
  s->mars_c -= s->mars_cd;
  if (s->mars_c < 0.0) s->mars_c += s->mars_cm;
  uni -= s->mars_c;
  if (uni < 0.0) uni += 1.0;
  return uni;
}
static int mars_init(ora_sim *s, int seed) {
  if (seed <= 0 || seed > 900000000) return -1; /* "Invalid seed for Marsaglia random # generator" */
  memset(s->mars_u, 0, sizeof(s->mars_u));
  int ij = (seed - 1) / 30082;
  int kl = (seed - 1) - 30082 * ij;
  int i = (ij / 177) % 177 + 2, j = ij % 177 + 2, k = (kl / 169) % 178 + 1, l = kl % 169;
  for (int ii = 1; ii <= 97; ii++) {
    double sum = 0.0, t = 0.5;
    for (int jj = 1; jj <= 24; jj++) {
      int m = ((i * j) % 179) * k % 179;
      i = j; j = k; k = m;
      l = (53 * l + 1) % 169;
      if ((l * m) % 64 >= 32) sum = sum + t;
      t = 0.5 * t;
    }
    s->mars_u[ii] = sum;
  }
  s->mars_c = 362436.0 / 16777216.0;
  s->mars_cd = 7654321.0 / 16777216.0;
  s->mars_cm = 16777213.0 / 16777216.0;
  s->mars_i97 = 97;
  s->mars_j97 = 33;
  mars_uniform(s);
  return 0;
}
/* the first n draws after construction (for checking the device generator) */
int ora_ranmars_stream(int seed, long n, double *out) {
  ora_sim *s = calloc(1, sizeof(ora_sim));
  if (mars_init(s, seed)) { free(s); return -1; }
  for (long k = 0; k < n; k++) out[k] = mars_uniform(s);
  free(s);
  return 0;
}

/* ---- fix langevin (fix_langevin.cpp): constructor :60-150 (seed + comm->me, ratio), init
 * :286-298 (gfactor1/2), setup :360-361 (post_force once per run set-up), compute_target
 * :773-787, post_force_templated<0,0,0,0,0,0> :576-713 — tally no, zero no, gjf no, per-type mass */
int ora_fix_langevin(ora_sim *s, double t_start, double t_stop, double t_period, int seed, const double *ratio, double dt) {
  if (t_period <= 0.0 || mars_init(s, seed)) return -1;
  s->langevin = 1;
  s->lang_tstart = t_start; s->lang_tstop = t_stop; s->lang_period = t_period; s->lang_dt = dt;
  s->lang_ratio = realloc(s->lang_ratio, sizeof(double) * (s->ntypes + 1));
  for (int t = 0; t <= s->ntypes; t++) s->lang_ratio[t] = ratio ? ratio[t] : 1.0;
  s->lang_begin = s->lang_end = s->step;
  return 0;
}
static void langevin_post_force(ora_sim *s) {
  double delta = s->step - s->lang_begin;
  if (delta != 0.0) delta /= s->lang_end - s->lang_begin;
  double t_target = s->lang_tstart + delta * (s->lang_tstop - s->lang_tstart);
  double tsqrt = sqrt(t_target);
  for (int i = 0; i < s->nlocal; i++) {
    int t = s->type[i];
    double gfactor1 = -s->mass[t] / s->lang_period / s->ftm2v;
    double gfactor2 = sqrt(s->mass[t]) / s->ftm2v;
    gfactor2 *= sqrt(24.0 * s->boltz / s->lang_period / s->lang_dt / s->mvv2e);
    gfactor1 *= 1.0 / s->lang_ratio[t];
    gfactor2 *= 1.0 / sqrt(s->lang_ratio[t]);
    double gamma1 = gfactor1, gamma2 = gfactor2 * tsqrt;
    double fran[3], fdrag[3];
    fran[0] = gamma2 * (mars_uniform(s) - 0.5);
    fran[1] = gamma2 * (mars_uniform(s) - 0.5);
    fran[2] = gamma2 * (mars_uniform(s) - 0.5);
    for (int d = 0; d < 3; d++) fdrag[d] = gamma1 * s->v[3 * i + d];
    for (int d = 0; d < 3; d++) s->f[3 * i + d] += fdrag[d] + fran[d];
  }
}

/* Verlet::setup (verlet.cpp:93-162); Modify::setup -> FixLangevin::setup applies the thermostat
 * to the set-up forces (delta = 0: Run::command sets beginstep = ntimestep before the set-up) */
int ora_setup(ora_sim *s) {
  if (ora_neigh_build(s)) return -1;
  if (ora_compute(s, 1, 1)) return -1;
  if (s->langevin) { s->lang_begin = s->step; s->lang_end = s->step; langevin_post_force(s); }
  return 0;
}

/* Verlet::run (verlet.cpp:229-360) with fix nve; thermo every `thermo_every` steps and at the end */
int ora_run(ora_sim *s, int nsteps, double dt, int thermo_every, ora_thermo *out, int max_out) {
  double dtv = dt, dtf = 0.5 * dt * s->ftm2v;
  long first = s->step, last = s->step + nsteps;
  int nout = 0;
  if (s->langevin) { s->lang_begin = first; s->lang_end = last; s->lang_dt = dt; }
  for (int i = 0; i < nsteps; i++) {
    s->step++;
    int out_step = (thermo_every > 0 && (s->step - first) % thermo_every == 0) || s->step == last;
    nve_initial(s, dtv, dtf);
    if (!decide(s)) forward_comm(s);
    else {
      s->in_run = 1;
      int rc = ora_neigh_build(s);
      s->in_run = 0;
      if (rc) return -1;
    }
    if (ora_compute(s, out_step, out_step)) return -1;
    if (s->langevin) langevin_post_force(s); /* Modify::post_force (verlet.cpp:337) */
    nve_final(s, dtf);
    if (out_step && out && nout < max_out) ora_thermo_compute(s, &out[nout++]);
  }
  return nout;
}

/* ---- accessors --------------------------------------------------------------------------- */
int ora_nlocal(ora_sim *s) { return s->nlocal; }
int ora_nghost(ora_sim *s) { return s->nghost; }
int ora_nbuilds(ora_sim *s) { return s->nbuilds; }
double *ora_x(ora_sim *s) { return s->x; }
double *ora_v(ora_sim *s) { return s->v; }
double *ora_f(ora_sim *s) { return s->f; }
int *ora_tag(ora_sim *s) { return s->tag; }
int *ora_type(ora_sim *s) { return s->type; }
int *ora_image(ora_sim *s) { return s->image; }
int *ora_numneigh(ora_sim *s) { return s->numneigh; }
long ora_neigh_total(ora_sim *s) { long n = 0; for (int i = 0; i < s->nlocal; i++) n += s->numneigh[i]; return n; }
/* neighbors as tags, atom after atom in local order */
void ora_neigh_tags(ora_sim *s, int *out) {
  long n = 0;
  for (int i = 0; i < s->nlocal; i++)
    for (int k = 0; k < s->numneigh[i]; k++) out[n++] = s->tag[s->neigh[s->firstoff[i] + k]];
}
void ora_bin_geometry(ora_sim *s, int *out12, double *out6) {
  int v[12] = {s->nbinx, s->nbiny, s->nbinz, s->mbinx, s->mbiny, s->mbinz, s->mbinxlo, s->mbinylo, s->mbinzlo, s->nstencil, 0, 0};
  memcpy(out12, v, sizeof(v));
  double d[6] = {s->binsizex, s->binsizey, s->binsizez, s->bininvx, s->bininvy, s->bininvz};
  memcpy(out6, d, sizeof(d));
}
