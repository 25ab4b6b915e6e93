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
#include <string.h>
#include <string>
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
