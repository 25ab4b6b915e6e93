// host.cpp — host-side geometry and decomposition logic of libb200md (no CUDA calls; usable
// and tested on a box without a GPU).
//
// Reference behaviour followed (/root/reference/src):
//   ProcMap::onelevel_grid / factor / best_factors   procmap.cpp:46-132, 742-762, 853-900
//   Domain::set_local_box                            domain.cpp:335-357
//   CommBrick::setup (SINGLE mode, orthogonal)       comm_brick.cpp:191-483
//   NBinStandard::setup_bins                         nbin_standard.cpp:82-214
//   NStencil::create_setup                           nstencil.cpp:213-247
//   CreateAtoms::loop_lattice / Lattice::lattice2box create_atoms.cpp:1301-1370, lattice.cpp:469-486
//   Velocity::create loop geom + RanPark             velocity.cpp:330-353, random_park.cpp:25-124
#include <math.h>
#include <stdint.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/b200md.h"

void b200md_set_error(const std::string &msg);

#define BIG 1.0e20
#define SMALL 1.0e-6
#define CUT2BIN_RATIO 100

extern "C" int b200md_procgrid(int nranks, const double prd[3], int procgrid[3]) {
  if (nranks < 1) { b200md_set_error("b200md: nranks must be >= 1"); return B200MD_ERR; }
  // all factorisations in the reference's enumeration order; first strictly-smaller surface wins
  const double area[3] = {prd[0] * prd[1], prd[0] * prd[2], prd[1] * prd[2]};
  double bestsurf = 2.0 * (area[0] + area[1] + area[2]);
  int best[3] = {nranks, 1, 1};
  bool found = false;
  for (int i = 1; i <= nranks; i++) {
    if (nranks % i) continue;
    const int nyz = nranks / i;
    for (int j = 1; j <= nyz; j++) {
      if (nyz % j) continue;
      const int k = nyz / j;
      const double surf = area[0] / i / j + area[1] / i / k + area[2] / j / k;
      if (surf < bestsurf) { bestsurf = surf; best[0] = i; best[1] = j; best[2] = k; found = true; }
    }
  }
  (void)found;
  procgrid[0] = best[0]; procgrid[1] = best[1]; procgrid[2] = best[2];
  return B200MD_OK;
}

extern "C" int b200md_rank2loc(int rank, const int procgrid[3], int myloc[3]) {
  const int n = procgrid[0] * procgrid[1] * procgrid[2];
  if (rank < 0 || rank >= n) { b200md_set_error("b200md: rank outside process grid"); return B200MD_ERR; }
  myloc[2] = rank % procgrid[2];
  myloc[1] = (rank / procgrid[2]) % procgrid[1];
  myloc[0] = rank / (procgrid[2] * procgrid[1]);
  return B200MD_OK;
}

static int loc2rank(const int loc[3], const int procgrid[3]) {
  return (loc[0] * procgrid[1] + loc[1]) * procgrid[2] + loc[2];
}

extern "C" int b200md_local_box(const double boxlo[3], const double boxhi[3], const int procgrid[3],
                                const int myloc[3], double sublo[3], double subhi[3]) {
  for (int d = 0; d < 3; d++) {
    const double prd = boxhi[d] - boxlo[d];
    // xsplit[i] = i * 1.0/procgrid (comm.cpp set_proc_grid)
    sublo[d] = boxlo[d] + prd * (myloc[d] * 1.0 / procgrid[d]);
    if (myloc[d] < procgrid[d] - 1) subhi[d] = boxlo[d] + prd * ((myloc[d] + 1) * 1.0 / procgrid[d]);
    else subhi[d] = boxhi[d];
  }
  return B200MD_OK;
}

extern "C" int b200md_comm_schedule(const double boxlo[3], const double boxhi[3], const int periodic[3],
                                    const int procgrid[3], const int myloc[3], double cutghost,
                                    b200md_swap_schedule *out) {
  memset(out, 0, sizeof(*out));
  double sublo[3], subhi[3], prd[3];
  b200md_local_box(boxlo, boxhi, procgrid, myloc, sublo, subhi);
  int maxneed[3], recvneed[3][2], sendneed[3][2], procneigh[3][2];
  for (int d = 0; d < 3; d++) {
    prd[d] = boxhi[d] - boxlo[d];
    maxneed[d] = static_cast<int>(cutghost * procgrid[d] / prd[d]) + 1;
    if (!periodic[d]) maxneed[d] = maxneed[d] < procgrid[d] - 1 ? maxneed[d] : procgrid[d] - 1;
    int loc[3] = {myloc[0], myloc[1], myloc[2]};
    int left = myloc[d] - 1; if (left < 0) left = procgrid[d] - 1;
    int right = myloc[d] + 1; if (right == procgrid[d]) right = 0;
    loc[d] = left; procneigh[d][0] = loc2rank(loc, procgrid);
    loc[d] = right; procneigh[d][1] = loc2rank(loc, procgrid);
    if (!periodic[d]) {
      auto mn = [](int a, int b) { return a < b ? a : b; };
      recvneed[d][0] = mn(maxneed[d], myloc[d]);
      recvneed[d][1] = mn(maxneed[d], procgrid[d] - myloc[d] - 1);
      sendneed[d][0] = mn(maxneed[d], procgrid[d] - left - 1);
      sendneed[d][1] = mn(maxneed[d], right);
    } else recvneed[d][0] = recvneed[d][1] = sendneed[d][0] = sendneed[d][1] = maxneed[d];
    out->maxneed[d] = maxneed[d];
  }
  (void)recvneed;
  int nswap = 2 * (maxneed[0] + maxneed[1] + maxneed[2]);
  if (nswap > B200MD_MAXSWAP) { b200md_set_error("b200md: too many swaps (box much smaller than cutoff)"); return B200MD_ERR; }
  out->nswap = nswap;
  int iswap = 0;
  for (int dim = 0; dim < 3; dim++)
    for (int ineed = 0; ineed < 2 * maxneed[dim]; ineed++, iswap++) {
      out->dim[iswap] = dim;
      out->ineed[iswap] = ineed;
      out->sendflag[iswap] = (ineed / 2 >= sendneed[dim][ineed % 2]) ? 0 : 1;
      if (ineed % 2 == 0) {
        out->sendproc[iswap] = procneigh[dim][0];
        out->recvproc[iswap] = procneigh[dim][1];
        out->slablo[iswap] = (ineed < 2) ? -BIG : 0.5 * (sublo[dim] + subhi[dim]);
        out->slabhi[iswap] = sublo[dim] + cutghost;
        if (myloc[dim] == 0) out->pbc[iswap][dim] = 1;
      } else {
        out->sendproc[iswap] = procneigh[dim][1];
        out->recvproc[iswap] = procneigh[dim][0];
        out->slablo[iswap] = subhi[dim] - cutghost;
        out->slabhi[iswap] = (ineed < 2) ? BIG : 0.5 * (sublo[dim] + subhi[dim]);
        if (myloc[dim] == procgrid[dim] - 1) out->pbc[iswap][dim] = -1;
      }
      // a non-periodic edge never shifts; its sendflag/sendneed already suppress the swap
      if (!periodic[dim]) out->pbc[iswap][dim] = 0;
    }
  return B200MD_OK;
}

// Direct halo (halo26.cu): the 26 directions (dx,dy,dz), index (dx+1) + 3(dy+1) + 9(dz+1), an owned atom can be a
// ghost in / migrate in.  For each: the rank a copy sent that way lands on (-1: no neighbour there, a non-periodic
// edge), the periodic shift multiples the sender applies (CommBrick::setup pbc flags, comm_brick.cpp:432-470), and
// per dimension the send slab bounds of the first swap (x <= slabhi_left sends left, x >= slablo_right sends right,
// comm_brick.cpp:407-428, 838).  usable = 0 when a dimension needs more than one ghost layer (maxneed > 1).
extern "C" int b200md_halo_directions(const double boxlo[3], const double boxhi[3], const int periodic[3],
                                      const int procgrid[3], const int myloc[3], double cutghost, int peer[27],
                                      int shift[27][3], double slabhi_left[3], double slablo_right[3], int *usable) {
  b200md_swap_schedule sch;
  if (b200md_comm_schedule(boxlo, boxhi, periodic, procgrid, myloc, cutghost, &sch)) return B200MD_ERR;
  *usable = (sch.maxneed[0] == 1 && sch.maxneed[1] == 1 && sch.maxneed[2] == 1) ? 1 : 0;
  for (int d = 0; d < 27; d++) { peer[d] = -1; shift[d][0] = shift[d][1] = shift[d][2] = 0; }
  if (!*usable) return B200MD_OK;
  for (int k = 0; k < 3; k++) { slabhi_left[k] = sch.slabhi[2 * k]; slablo_right[k] = sch.slablo[2 * k + 1]; }
  for (int d = 0; d < 27; d++) {
    if (d == 13) continue;
    const int comp[3] = {d % 3 - 1, (d / 3) % 3 - 1, d / 9 - 1};
    int loc[3] = {myloc[0], myloc[1], myloc[2]};
    bool ok = true;
    for (int k = 0; k < 3 && ok; k++) {
      if (!comp[k]) continue;
      const int sw = 2 * k + (comp[k] < 0 ? 0 : 1);   // maxneed == 1: swap 2k sends left, 2k+1 right
      if (!sch.sendflag[sw]) { ok = false; break; }
      loc[k] += comp[k];
      if (loc[k] < 0) loc[k] = procgrid[k] - 1;
      if (loc[k] >= procgrid[k]) loc[k] = 0;
      shift[d][k] = sch.pbc[sw][k];
    }
    if (ok) peer[d] = loc2rank(loc, procgrid);
    else shift[d][0] = shift[d][1] = shift[d][2] = 0;
  }
  return B200MD_OK;
}

extern "C" int b200md_bin_setup(const double boxlo[3], const double boxhi[3], const double sublo[3],
                                const double subhi[3], double cutghost, double cutneighmax,
                                b200md_bin_geometry *g) {
  double bbox[3], bsublo[3], bsubhi[3];
  for (int d = 0; d < 3; d++) {
    bsublo[d] = sublo[d] - cutghost;
    bsubhi[d] = subhi[d] + cutghost;
    bbox[d] = boxhi[d] - boxlo[d];
  }
  double binsize_optimal = 0.5 * cutneighmax;
  if (binsize_optimal == 0.0) binsize_optimal = bbox[0];
  const double binsizeinv = 1.0 / binsize_optimal;
  if (bbox[0] * binsizeinv > 2147483647.0 || bbox[1] * binsizeinv > 2147483647.0 || bbox[2] * binsizeinv > 2147483647.0) {
    b200md_set_error("Domain too large for neighbor bins");
    return B200MD_ERR;
  }
  int nbin[3];
  double binsize[3], bininv[3];
  for (int d = 0; d < 3; d++) {
    nbin[d] = static_cast<int>(bbox[d] * binsizeinv);
    if (nbin[d] == 0) nbin[d] = 1;
    binsize[d] = bbox[d] / nbin[d];
    bininv[d] = 1.0 / binsize[d];
    if (binsize_optimal * bininv[d] > CUT2BIN_RATIO) {
      b200md_set_error("Cannot use neighbor bins - box size << cutoff");
      return B200MD_ERR;
    }
  }
  int mlo[3], mhi[3];
  for (int d = 0; d < 3; d++) {
    double coord = bsublo[d] - SMALL * bbox[d];
    mlo[d] = static_cast<int>((coord - boxlo[d]) * bininv[d]);
    if (coord < boxlo[d]) mlo[d] = mlo[d] - 1;
    coord = bsubhi[d] + SMALL * bbox[d];
    mhi[d] = static_cast<int>((coord - boxlo[d]) * bininv[d]);
    mlo[d] -= 1;
    mhi[d] += 1;
  }
  g->nbinx = nbin[0]; g->nbiny = nbin[1]; g->nbinz = nbin[2];
  g->mbinxlo = mlo[0]; g->mbinylo = mlo[1]; g->mbinzlo = mlo[2];
  g->mbinx = mhi[0] - mlo[0] + 1; g->mbiny = mhi[1] - mlo[1] + 1; g->mbinz = mhi[2] - mlo[2] + 1;
  g->binsizex = binsize[0]; g->binsizey = binsize[1]; g->binsizez = binsize[2];
  g->bininvx = bininv[0]; g->bininvy = bininv[1]; g->bininvz = bininv[2];
  // NStencil::create_setup
  g->sx = static_cast<int>(cutneighmax * bininv[0]); if (g->sx * binsize[0] < cutneighmax) g->sx++;
  g->sy = static_cast<int>(cutneighmax * bininv[1]); if (g->sy * binsize[1] < cutneighmax) g->sy++;
  g->sz = static_cast<int>(cutneighmax * bininv[2]); if (g->sz * binsize[2] < cutneighmax) g->sz++;
  return B200MD_OK;
}

// ---- RanPark (random_park.cpp) ----
namespace {
struct RanPark {
  int seed;
  double uniform() {
    const int IA = 16807, IM = 2147483647, IQ = 127773, IR = 2836;
    const double AM = 1.0 / IM;
    int k = seed / IQ;
    seed = IA * (seed - k * IQ) - IR * k;
    if (seed < 0) seed += IM;
    return AM * seed;
  }
  void reset(int ibase, const double *coord) {
    // Jenkins one-at-a-time hash over the bytes of the seed and the three coordinates;
    // bytes are the platform's (signed) char like the reference
    unsigned int hash = 0;
    const char *str = reinterpret_cast<const char *>(&ibase);
    for (int i = 0; i < (int)sizeof(int); i++) { hash += str[i]; hash += (hash << 10); hash ^= (hash >> 6); }
    str = reinterpret_cast<const char *>(coord);
    for (int i = 0; i < (int)(3 * sizeof(double)); i++) { hash += str[i]; hash += (hash << 10); hash ^= (hash >> 6); }
    hash += (hash << 3);
    hash ^= (hash >> 11);
    hash += (hash << 15);
    seed = hash & 0x7ffffff;
    if (!seed) seed = 1;
    for (int i = 0; i < 5; i++) uniform();
  }
};
}  // namespace

extern "C" int b200md_lattice_fcc(double scale, int nx, int ny, int nz, const double sublo[3],
                                  const double subhi[3], int seed, double mass, int *natoms_out,
                                  double *x_aos, double *v_aos, int *tag) {
  if (seed <= 0) { b200md_set_error("Illegal velocity create seed argument"); return B200MD_ERR; }
  static const double basis[4][3] = {{0, 0, 0}, {0.5, 0.5, 0}, {0.5, 0, 0.5}, {0, 0.5, 0.5}};
  const double factor = 1.0 / sqrt(mass);
  RanPark rng;
  rng.seed = 1;
  long long n = 0;
  for (int k = 0; k < nz; k++)
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++)
        for (int m = 0; m < 4; m++) {
          // lattice2box with identity orientation: primitive rows (1,0,0)… then *scale
          double x[3] = {i + basis[m][0], j + basis[m][1], k + basis[m][2]};
          x[0] = x[0] * 1.0 + x[1] * 0.0 + x[2] * 0.0;  // kept for operand order parity
          x[0] *= scale; x[1] *= scale; x[2] *= scale;
          if (x[0] < sublo[0] || x[0] >= subhi[0] || x[1] < sublo[1] || x[1] >= subhi[1] ||
              x[2] < sublo[2] || x[2] >= subhi[2]) continue;
          if (x_aos) {
            x_aos[3 * n] = x[0]; x_aos[3 * n + 1] = x[1]; x_aos[3 * n + 2] = x[2];
            if (tag) tag[n] = (int)((((long long)k * ny + j) * nx + i) * 4 + m + 1);
            if (v_aos) {
              rng.reset(seed, x);
              const double vx = rng.uniform() - 0.5, vy = rng.uniform() - 0.5, vz = rng.uniform() - 0.5;
              v_aos[3 * n] = vx * factor; v_aos[3 * n + 1] = vy * factor; v_aos[3 * n + 2] = vz * factor;
            }
          }
          n++;
        }
  if (n > 2147483647LL) { b200md_set_error("b200md: too many atoms for one rank"); return B200MD_ERR; }
  *natoms_out = (int)n;
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// Atom::sort — the local index order the reference's set-up leaves (atom.cpp:2246-2349 with
// setup_sort_bins :2351-2410): bins of cutneighmax/2 over the (sub-)box, stable inside a bin.
// The device keeps its own bin order; this order only matters to callers that must reproduce
// order-dependent host semantics (fix langevin's RanMars draws, fix_langevin.cpp:636-662).
// ---------------------------------------------------------------------------------------
extern "C" int b200md_atom_sort_order(int n, const double *x, const double sublo[3], const double subhi[3],
                                      double cutneighmax, int *permute) {
  if (n < 0 || (n && (!x || !permute)) || !sublo || !subhi) { b200md_set_error("b200md: null argument"); return B200MD_ERR; }
  const double binsize = 0.5 * cutneighmax;
  int nb[3] = {1, 1, 1};
  double binv[3] = {0, 0, 0};
  if (binsize > 0.0) {
    const double bininv = 1.0 / binsize;
    for (int d = 0; d < 3; d++) {
      nb[d] = static_cast<int>((subhi[d] - sublo[d]) * bininv);
      if (nb[d] == 0) nb[d] = 1;
      binv[d] = nb[d] / (subhi[d] - sublo[d]);
    }
  }
  const long long nbins = (long long)nb[0] * nb[1] * nb[2];
  if (binsize <= 0.0 || nbins == 1) {   // sorting disabled / single bin: order unchanged
    for (int i = 0; i < n; i++) permute[i] = i;
    return B200MD_OK;
  }
  std::vector<int> count(nbins + 1, 0), bin(n);
  for (int i = 0; i < n; i++) {
    int c[3];
    for (int d = 0; d < 3; d++) {
      c[d] = static_cast<int>((x[3 * (size_t)i + d] - sublo[d]) * binv[d]);
      c[d] = c[d] < 0 ? 0 : c[d];
      c[d] = c[d] > nb[d] - 1 ? nb[d] - 1 : c[d];
    }
    bin[i] = (c[2] * nb[1] + c[1]) * nb[0] + c[0];
    count[bin[i] + 1]++;
  }
  for (long long b = 0; b < nbins; b++) count[b + 1] += count[b];
  for (int i = 0; i < n; i++) permute[count[bin[i]]++] = i;   // new atom I = old atom permute[I]
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// RanMars in integer form (random_mars.cpp:30-95).  Every value the generator holds is a
// multiple of 2^-24, so with L the lagged-Fibonacci sequence in units of 2^-24
//     L[n+97] = L[n] - L[n+64]            (mod 2^24)         u[i97] - u[j97], lags 97 and 33
//     c[n]    = (362436 - (n+1) 7654321)  (mod 16777213)     the arithmetic sequence `c`
//     draw n  = ((L[n+97] - c[n]) mod 2^24) 2^-24            n = 0 is the constructor's own draw
// The recurrence is linear, so the state k draws ahead is a polynomial in the shift operator:
// x^k mod (x^97 + x^64 - 1) over Z/2^24 — computed here by square-and-multiply, applied on the
// device by every block of the stream kernel to reach its own chunk (langevin.cu).
// ---------------------------------------------------------------------------------------
namespace {
const int LAG = B200MD_RANMARS_LAG;
typedef uint32_t u32;

void poly_reduce(u32 *c, int top) {   // x^97 = 1 - x^64
  for (int d = top; d >= LAG; d--) {
    const u32 a = c[d];
    if (!a) continue;
    c[d] = 0;
    c[d - LAG] += a;
    c[d - 33] -= a;
  }
}
void poly_mul(const u32 *a, const u32 *b, u32 *out) {
  u32 c[2 * LAG];
  memset(c, 0, sizeof(c));
  for (int i = 0; i < LAG; i++) {
    if (!a[i]) continue;
    for (int j = 0; j < LAG; j++) c[i + j] += a[i] * b[j];
  }
  poly_reduce(c, 2 * LAG - 2);
  memcpy(out, c, sizeof(u32) * LAG);
}
void poly_pow(long long k, u32 *out) {   // x^k mod P
  u32 result[LAG], base[LAG];
  memset(result, 0, sizeof(result)); memset(base, 0, sizeof(base));
  result[0] = 1; base[1] = 1;
  while (k > 0) {
    if (k & 1) poly_mul(result, base, result);
    k >>= 1;
    if (k) poly_mul(base, base, base);
  }
  memcpy(out, result, sizeof(result));
}
}  // namespace

// rows i = 0..96 of the jump-by-k matrix: coefficients of x^(k+i) mod P, so that
// L[p+k+i] = sum_j M[i][j] L[p+j]  (mod 2^24; the arithmetic may run mod 2^32 and be masked at the end)
void ranmars_jump_matrix(long long k, uint32_t *M) {
  u32 row[LAG + 1];
  poly_pow(k, row);
  for (int i = 0; i < LAG; i++) {
    memcpy(M + (size_t)i * LAG, row, sizeof(u32) * LAG);
    // multiply by x
    const u32 top = row[LAG - 1];
    for (int j = LAG - 1; j > 0; j--) row[j] = row[j - 1];
    row[0] = top;
    row[64] -= top;
  }
}

extern "C" int b200md_ranmars_state(int seed, uint32_t state[B200MD_RANMARS_LAG]) {
  if (seed <= 0 || seed > 900000000) { b200md_set_error("Invalid seed for Marsaglia random # generator"); return B200MD_ERR; }
  u32 u[LAG + 1];
  int ij = (seed - 1) / 30082;
  int kl = (seed - 1) - 30082 * ij;
  int i = (ij / 177) % 177 + 2, j = ij % 177 + 2, k = (kl / 169) % 178 + 1, l = kl % 169;
  for (int ii = 1; ii <= LAG; ii++) {
    u32 s = 0;
    for (int jj = 1; jj <= 24; jj++) {
      const int m = ((i * j) % 179) * k % 179;
      i = j; j = k; k = m;
      l = (53 * l + 1) % 169;
      if ((l * m) % 64 >= 32) s |= 1u << (24 - jj);   // s += 2^-jj
    }
    u[ii] = s;
  }
  // call n reads slot 97-n (n < 97): L[n] = u[97-n]; the constructor's draw moves the window to position 1
  u32 L[LAG + 1];
  for (int n = 0; n < LAG; n++) L[n] = u[LAG - n];
  L[LAG] = (L[0] - L[64]) & 0xffffffu;
  for (int n = 0; n < LAG; n++) state[n] = L[n + 1];
  return B200MD_OK;
}

extern "C" int b200md_ranmars_jump(const uint32_t state[B200MD_RANMARS_LAG], long long k, uint32_t out[B200MD_RANMARS_LAG]) {
  if (!state || !out || k < 0) { b200md_set_error("b200md: bad ranmars_jump arguments"); return B200MD_ERR; }
  std::vector<u32> M((size_t)LAG * LAG);
  ranmars_jump_matrix(k, M.data());
  u32 r[LAG];
  for (int i = 0; i < LAG; i++) {
    u32 acc = 0;
    for (int j = 0; j < LAG; j++) acc += M[(size_t)i * LAG + j] * state[j];
    r[i] = acc & 0xffffffu;
  }
  memcpy(out, r, sizeof(r));
  return B200MD_OK;
}

extern "C" int b200md_ranmars_draws(const uint32_t state[B200MD_RANMARS_LAG], long long position, int count, double *out) {
  if (!state || position < 0 || count < 0 || (count && !out)) { b200md_set_error("b200md: bad ranmars_draws arguments"); return B200MD_ERR; }
  std::vector<u32> L(state, state + LAG);
  L.resize((size_t)LAG + count);
  for (int k = 0; k < count; k++) {
    const u32 w = (L[k] - L[k + 64]) & 0xffffffu;
    L[(size_t)LAG + k] = w;
    const long long n = position + k;
    const long long cc = (362436LL + 16777213LL - ((((n + 1) % 16777213LL) * 7654321LL) % 16777213LL)) % 16777213LL;
    out[k] = (double)((w - (u32)cc) & 0xffffffu) * (1.0 / 16777216.0);
  }
  return B200MD_OK;
}
