// neigh.cu — binning, spatial sort and neighbor-list construction on the device.
//
// Follows the behaviour of (reference, /root/reference/src):
//   NBin::coord2bin                    nbin.cpp:139-168        -> coord2bin()
//   NBinStandard::bin_atoms            nbin_standard.cpp:221-260 -> k_bin_count/k_bin_scatter/k_cell_sort
//   NStencil{Full,Half}Bin3d::create   nstencil_full_bin_3d.cpp:26-37, nstencil_half_bin_3d.cpp:26-38
//                                       -> stencil rows (neigh_setup_bins)
//   NPairFullBinAtomonly::build        npair_full_bin_atomonly.cpp:32-92   -> k_build<…,false>
//   NPairHalfBinAtomonlyNewton::build  npair_half_bin_atomonly_newton.cpp:33-113 -> k_build<…,true>
//   Domain::pbc                        domain.cpp:530-650      -> k_pbc
//   Atom::sort                         atom.cpp:2246-2349      -> neigh_sort_owned (every rebuild)
// Design differences (B200-first): cells are contiguous ranges of a counting sort instead of
// linked lists; a stencil is a set of (dy,dz) rows, each one contiguous range of the
// cell-ordered position copy `cell_xq`, so candidate reads are coalesced 32-byte records;
// TPA lanes cooperate on one atom and compact hits with a ballot; the list is written
// straight into the warp-tile-transposed layout (ctx.h).  One pass, fixed pitch; overflow
// regrows and reruns.
#include <stdio.h>
#include <algorithm>
#include "ctx.h"
#include "util.cuh"

struct BinParams {
  double bboxlo[3], bboxhi[3];
  double bininv[3];
  int nbin[3], mbinlo[3], mbin[3];
};

static BinParams make_bin_params(const b200md_ctx *c) {
  BinParams p;
  for (int d = 0; d < 3; d++) { p.bboxlo[d] = c->boxlo[d]; p.bboxhi[d] = c->boxhi[d]; }
  p.bininv[0] = c->bin.bininvx; p.bininv[1] = c->bin.bininvy; p.bininv[2] = c->bin.bininvz;
  p.nbin[0] = c->bin.nbinx; p.nbin[1] = c->bin.nbiny; p.nbin[2] = c->bin.nbinz;
  p.mbinlo[0] = c->bin.mbinxlo; p.mbinlo[1] = c->bin.mbinylo; p.mbinlo[2] = c->bin.mbinzlo;
  p.mbin[0] = c->bin.mbinx; p.mbin[1] = c->bin.mbiny; p.mbin[2] = c->bin.mbinz;
  return p;
}

__device__ __forceinline__ int coord2bin1(double x, double lo, double hi, double inv, int nbin) {
  // three branches exactly as nbin.cpp:146-152; products rounded separately (no FMA)
  if (x >= hi) return (int)(__dmul_rn(__dsub_rn(x, hi), inv)) + nbin;
  if (x >= lo) {
    int i = (int)(__dmul_rn(__dsub_rn(x, lo), inv));
    return min(i, nbin - 1);
  }
  return (int)(__dmul_rn(__dsub_rn(x, lo), inv)) - 1;
}

__device__ __forceinline__ int coord2bin(const double4 &q, const BinParams &p, bool &ok) {
  int ix = coord2bin1(q.x, p.bboxlo[0], p.bboxhi[0], p.bininv[0], p.nbin[0]) - p.mbinlo[0];
  int iy = coord2bin1(q.y, p.bboxlo[1], p.bboxhi[1], p.bininv[1], p.nbin[1]) - p.mbinlo[1];
  int iz = coord2bin1(q.z, p.bboxlo[2], p.bboxhi[2], p.bininv[2], p.nbin[2]) - p.mbinlo[2];
  ok = isfinite(q.x) && isfinite(q.y) && isfinite(q.z) && ix >= 0 && ix < p.mbin[0] &&
       iy >= 0 && iy < p.mbin[1] && iz >= 0 && iz < p.mbin[2];
  return (iz * p.mbin[1] + iy) * p.mbin[0] + ix;
}

// ---------------------------------------------------------------------------------------
// per-atom array growth
// ---------------------------------------------------------------------------------------
template <typename T>
static int regrow(T *&p, size_t oldn, size_t newn, cudaStream_t s) {
  T *q = nullptr;
  CUDA_TRY(cudaMalloc(&q, newn * sizeof(T)));
  if (p && oldn) CUDA_TRY(cudaMemcpyAsync(q, p, oldn * sizeof(T), cudaMemcpyDeviceToDevice, s));
  if (p) { CUDA_TRY(cudaStreamSynchronize(s)); CUDA_TRY(cudaFree(p)); }
  p = q;
  return B200MD_OK;
}
// SoA 3-vector with stride = capacity: regrow keeps component layout
static int regrow3(double *&p, size_t oldcap, size_t newcap, size_t used, cudaStream_t s) {
  double *q = nullptr;
  CUDA_TRY(cudaMalloc(&q, 3 * newcap * sizeof(double)));
  CUDA_TRY(cudaMemsetAsync(q, 0, 3 * newcap * sizeof(double), s));
  if (p && used)
    for (int d = 0; d < 3; d++)
      CUDA_TRY(cudaMemcpyAsync(q + d * newcap, p + d * oldcap, used * sizeof(double),
                               cudaMemcpyDeviceToDevice, s));
  if (p) { CUDA_TRY(cudaStreamSynchronize(s)); CUDA_TRY(cudaFree(p)); }
  p = q;
  return B200MD_OK;
}

int neigh_grow_atoms(b200md_ctx *c, int need) {
  if (need <= c->nmax) return B200MD_OK;
  size_t oldcap = c->nmax;
  size_t newcap = (size_t)(need * 1.15) + 1024;
  newcap = (newcap + 31) & ~(size_t)31;
  size_t used = c->nlocal + c->nghost;
  if (used > oldcap) used = oldcap;
  cudaStream_t s = c->stream;
  TRY(regrow(c->xq, used, newcap, s));
  TRY(regrow(c->xq_tmp, 0, newcap, s));
  TRY(regrow(c->cell_xq, 0, newcap, s));
  TRY(regrow3(c->v, oldcap, newcap, used, s));
  TRY(regrow3(c->v_tmp, oldcap, newcap, 0, s));
  TRY(regrow3(c->f, oldcap, newcap, used, s));
  TRY(regrow3(c->xhold, oldcap, newcap, used, s));
  TRY(regrow(c->tag, used, newcap, s));
  TRY(regrow(c->tag_tmp, 0, newcap, s));
  TRY(regrow(c->image, used, newcap, s));
  TRY(regrow(c->image_tmp, 0, newcap, s));
  TRY(regrow(c->atom_bin, 0, newcap, s));
  TRY(regrow(c->cell_atoms, 0, newcap, s));
  TRY(regrow(c->perm, 0, newcap, s));
  TRY(regrow(c->numneigh, 0, newcap, s));
  TRY(regrow(c->ghost_src, used, newcap, s));
  TRY(regrow(c->ghost_pbc, used, newcap, s));
  TRY(regrow(c->ghost_rank, used, newcap, s));
  TRY(regrow(c->rho, 0, newcap, s));
  TRY(regrow(c->fp, 0, newcap, s));
  c->nmax = (int)newcap;
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// bin geometry + stencil rows (host)
// ---------------------------------------------------------------------------------------
static double bin_distance(const b200md_bin_geometry &g, int i, int j, int k) {
  // nstencil.cpp:405-422
  double delx, dely, delz;
  if (i > 0) delx = (i - 1) * g.binsizex; else if (i == 0) delx = 0.0; else delx = (i + 1) * g.binsizex;
  if (j > 0) dely = (j - 1) * g.binsizey; else if (j == 0) dely = 0.0; else dely = (j + 1) * g.binsizey;
  if (k > 0) delz = (k - 1) * g.binsizez; else if (k == 0) delz = 0.0; else delz = (k + 1) * g.binsizez;
  return delx * delx + dely * dely + delz * delz;
}

// Bin size and list layout for the system as it is configured now (pair style and topology may arrive after
// b200md_neigh_config): called at the start of every rebuild, re-derives the bins only when the choice changed.
// Measured on B200 at 2.05 M atoms: cluster rows want about one 4-atom cluster per bin (1.2 x the
// reference's cutneighmax/2: the cluster is then roughly cubic, pair 0.436 -> 0.402 ms), per-atom rows of atomic systems
// thin bin rows (0.6 x: the lists of the 32 atoms of a warp become shifted copies of one another, eam pair 1.13 -> 0.99 ms),
// short-cutoff molecular systems fewer bins (2 x: in.chain has 5.5 M bins to scan at 1 x).
int neigh_choose_layout(b200md_ctx *c) {
  const bool cluster = !c->topo.molecular && (c->use_cluster < 0 ? c->pair_style == B200MD_PAIR_LJ_CUT : c->use_cluster != 0);
  const double want = c->bin_scale_user > 0 ? c->bin_scale_user : (cluster ? 1.2 : (c->topo.molecular ? 2.0 : 0.6));
  if (want != c->bin_scale) {
    c->bin_scale = want;
    TRY(neigh_setup_bins(c));
  }
  return B200MD_OK;
}

int neigh_setup_bins(b200md_ctx *c) {
  // bins of bin_scale x cutneighmax/2 (1 = the reference's size, NBinStandard::setup_bins).  List membership is decided by
  // distance, so the bin size is a device-side choice: the stencil rows below are cut by bin_distance < cutneighmax^2
  // whatever extents the geometry reports.
  TRY(b200md_bin_setup(c->boxlo, c->boxhi, c->sublo, c->subhi, c->cutghost, c->cutneighmax * c->bin_scale, &c->bin));
  if (c->bin_scale != 1.0) {   // stencil extents for the real cutoff (NStencil::create_setup, nstencil.cpp:213-247)
    b200md_bin_geometry &b = c->bin;
    b.sx = (int)(c->cutneighmax * b.bininvx); if (b.sx * b.binsizex < c->cutneighmax) b.sx++;
    b.sy = (int)(c->cutneighmax * b.bininvy); if (b.sy * b.binsizey < c->cutneighmax) b.sy++;
    b.sz = (int)(c->cutneighmax * b.bininvz); if (b.sz * b.binsizez < c->cutneighmax) b.sz++;
  }
  const b200md_bin_geometry &g = c->bin;
  long long mb = (long long)g.mbinx * g.mbiny * g.mbinz;
  if (mb + 1 > 2147483647LL) { b200md_set_error("Too many neighbor bins"); return B200MD_ERR; }
  c->mbins = (int)mb;
  if (c->mbins + 1 > c->cell_cap) {
    if (c->cell_count) { cudaFree(c->cell_count); cudaFree(c->cell_start); cudaFree(c->cell_fill); }
    c->cell_cap = c->mbins + 1 + 1024;
    CUDA_TRY(cudaMalloc(&c->cell_count, (size_t)c->cell_cap * sizeof(int)));
    CUDA_TRY(cudaMalloc(&c->cell_start, (size_t)c->cell_cap * sizeof(int)));
    CUDA_TRY(cudaMalloc(&c->cell_fill, (size_t)c->cell_cap * sizeof(int)));
  }
  // stencil rows: for each (dy,dz) the included dx form [-sxr,sxr] (bin_distance is
  // monotone in |dx|); membership test `bin_distance < cutneighmaxsq`
  // (nstencil_full_bin_3d.cpp:35).  Half stencil (nstencil_half_bin_3d.cpp:33-36) keeps
  // k>0 || j>0 || (j==0 && i>0): rows with dz>0 or (dz==0,dy>0) in full, the own row only
  // for dx>0; the own bin gets the in-bin ordering rule.
  const double cutsq = c->cutneighmax * c->cutneighmax;
  std::vector<int> off, sxr, flag;
  for (int k = -g.sz; k <= g.sz; k++)
    for (int j = -g.sy; j <= g.sy; j++) {
      int ext = -1;
      for (int i = 0; i <= g.sx; i++)
        if (bin_distance(g, i, j, k) < cutsq) ext = i;
      if (ext < 0) continue;
      int role = 1;
      if (c->list_kind == B200MD_LIST_HALF_NEWTON) {
        if (k > 0 || (k == 0 && j > 0)) role = 1;
        else if (k == 0 && j == 0) role = 2;
        else continue;
      }
      off.push_back((k * g.mbiny + j) * g.mbinx);
      sxr.push_back(ext);
      flag.push_back(role);
    }
  c->nrows = (int)off.size();
  if (c->row_off) { cudaFree(c->row_off); cudaFree(c->row_sx); cudaFree(c->row_flag); }
  CUDA_TRY(cudaMalloc(&c->row_off, sizeof(int) * std::max(1, c->nrows)));
  CUDA_TRY(cudaMalloc(&c->row_sx, sizeof(int) * std::max(1, c->nrows)));
  CUDA_TRY(cudaMalloc(&c->row_flag, sizeof(int) * std::max(1, c->nrows)));
  CUDA_TRY(cudaMemcpy(c->row_off, off.data(), sizeof(int) * c->nrows, cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->row_sx, sxr.data(), sizeof(int) * c->nrows, cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(c->row_flag, flag.data(), sizeof(int) * c->nrows, cudaMemcpyHostToDevice));
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// exclusive scan of int array (three small kernels; n ~ 1e6 bins)
// ---------------------------------------------------------------------------------------
#define SCAN_T 256
#define SCAN_ITEMS 8
#define SCAN_TILE (SCAN_T * SCAN_ITEMS)

__device__ __forceinline__ int block_excl_scan(int v, int *s_warp, int &total) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) s_warp[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    int w = lane < (SCAN_T / 32) ? s_warp[lane] : 0;
    int winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += t;
    }
    if (lane < (SCAN_T / 32)) s_warp[lane] = winc - w;
    if (lane == 31) s_warp[32] = winc;
  }
  __syncthreads();
  total = s_warp[32];
  int r = inc - v + s_warp[wid];
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(SCAN_T) k_scan_tiles(const int *__restrict__ in, int *__restrict__ out,
                                                       int *__restrict__ tile_sum, int n) {
  __shared__ int s_warp[33];
  const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS], sum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { v[k] = (base + k < n) ? in[base + k] : 0; sum += v[k]; }
  int total, ex = block_excl_scan(sum, s_warp, total);
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) { if (base + k < n) out[base + k] = ex; ex += v[k]; }
  if (threadIdx.x == 0) tile_sum[blockIdx.x] = total;
}
__global__ void __launch_bounds__(SCAN_T) k_scan_sums(int *tile_sum, int ntiles) {
  // single block; ntiles <= SCAN_TILE * something: loop with carry
  __shared__ int s_warp[33];
  int carry = 0;
  for (int base = 0; base < ntiles; base += SCAN_T) {
    int i = base + threadIdx.x;
    int v = i < ntiles ? tile_sum[i] : 0;
    int total, ex = block_excl_scan(v, s_warp, total);
    if (i < ntiles) tile_sum[i] = ex + carry;
    carry += total;
  }
}
__global__ void __launch_bounds__(SCAN_T) k_scan_add(int *__restrict__ out, const int *__restrict__ tile_sum, int n) {
  const int add = tile_sum[blockIdx.x];
  const int base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; k++) if (base + k < n) out[base + k] += add;
}

int dev_exclusive_scan(b200md_ctx *c, const int *in, int *out, int n) {
  int ntiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  if (ntiles > c->scan_cap) {
    if (c->scan_block) cudaFree(c->scan_block);
    c->scan_cap = ntiles + 64;
    CUDA_TRY(cudaMalloc(&c->scan_block, sizeof(int) * c->scan_cap));
  }
  k_scan_tiles<<<ntiles, SCAN_T, 0, c->stream>>>(in, out, c->scan_block, n);
  k_scan_sums<<<1, SCAN_T, 0, c->stream>>>(c->scan_block, ntiles);
  k_scan_add<<<ntiles, SCAN_T, 0, c->stream>>>(out, c->scan_block, n);
  c->launches += 3;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// pbc, binning, sort
// ---------------------------------------------------------------------------------------
#define IMGMASK 1023
#define IMGBITS 10
#define IMG2BITS 20

struct PbcParams { double lo[3], hi[3], prd[3]; int periodic[3]; };

__global__ void k_pbc(double4 *__restrict__ xq, int *__restrict__ image, int n, PbcParams p, int *flags) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 q = xq[i];
  if (!(isfinite(q.x) && isfinite(q.y) && isfinite(q.z))) { flags[2] = 1; return; }
  int img = image[i];
  double xc[3] = {q.x, q.y, q.z};
  bool changed = false;
#pragma unroll
  for (int d = 0; d < 3; d++) {
    if (!p.periodic[d]) continue;
    const int sh = d * IMGBITS;
    if (xc[d] < p.lo[d]) {
      xc[d] = __dadd_rn(xc[d], p.prd[d]);
      int idim = (img >> sh) & IMGMASK;
      int other = img ^ (idim << sh);
      idim = (idim - 1) & IMGMASK;
      img = other | (idim << sh);
      changed = true;
    }
    if (xc[d] >= p.hi[d]) {
      xc[d] = __dsub_rn(xc[d], p.prd[d]);
      xc[d] = fmax(xc[d], p.lo[d]);
      int idim = (img >> sh) & IMGMASK;
      int other = img ^ (idim << sh);
      idim = (idim + 1) & IMGMASK;
      img = other | (idim << sh);
      changed = true;
    }
  }
  if (changed) {
    q.x = xc[0]; q.y = xc[1]; q.z = xc[2];
    xq[i] = q;
    image[i] = img;
  }
}

int neigh_pbc(b200md_ctx *c) {
  if (c->nlocal == 0) return B200MD_OK;
  PbcParams p;
  for (int d = 0; d < 3; d++) { p.lo[d] = c->boxlo[d]; p.hi[d] = c->boxhi[d]; p.prd[d] = c->prd[d]; p.periodic[d] = c->periodic[d]; }
  k_pbc<<<(c->nlocal + 255) / 256, 256, 0, c->stream>>>(c->xq, c->image, c->nlocal, p, c->d_flags);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

__global__ void k_bin_count(const double4 *__restrict__ xq, int n, BinParams p, int *__restrict__ atom_bin,
                            int *__restrict__ cell_count, int *flags, const int *__restrict__ dead, int dead_bin) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  bool ok;
  int b = coord2bin(xq[i], p, ok);
  if (dead && dead[i]) b = dead_bin;   // migrated away (halo26.cu): parked behind the last bin, dropped by the permutation
  else if (!ok) { flags[2] = 1; b = 0; }
  atom_bin[i] = b;
  atomicAdd(&cell_count[b], 1);
}
__global__ void k_bin_scatter(const int *__restrict__ atom_bin, int n, const int *__restrict__ cell_start,
                              int *__restrict__ cell_fill, int *__restrict__ cell_atoms) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int b = atom_bin[i];
  int pos = cell_start[b] + atomicAdd(&cell_fill[b], 1);
  cell_atoms[pos] = i;
}
// make the within-cell order deterministic.  BYTAG = false: ascending atom index = the order of the reference's
// linked lists (nbin_standard.cpp:252-257: owned ascending, ghosts last).  BYTAG = true (direct halo on several
// ranks, where arrivals and ghosts are placed through atomic slot counters): owned atoms by tag, then ghosts by
// (tag, z, y, x) — independent of the order the messages were filled in, and of the decomposition.
// BYX (option cell_order x): ascending x first.  Atoms that follow each other in memory then follow each other along
// the bin row, so the lists of the 32 atoms of a warp are shifted copies of one another and their gathers and REDs of
// the same iteration fall into the same lines (eam, per-atom rows: pair time -5 %).
template <bool BYTAG, bool BYX>
__device__ __forceinline__ bool cell_before(int a, int b, int nlocal, const int *__restrict__ tag, const double4 *__restrict__ xq) {
  if (BYX) {
    const double xa = xq[a].x, xb = xq[b].x;
    if (xa != xb) return xa < xb;
  }
  if (!BYTAG) return a < b;
  const bool ga = a >= nlocal, gb = b >= nlocal;
  if (ga != gb) return gb;
  const int ta = tag[a], tb = tag[b];
  if (ta != tb) return ta < tb;
  if (!ga) return a < b;
  const double4 qa = xq[a], qb = xq[b];
  if (qa.z != qb.z) return qa.z < qb.z;
  if (qa.y != qb.y) return qa.y < qb.y;
  return qa.x < qb.x;
}
template <bool BYTAG, bool BYX>
__global__ void k_cell_sort(const int *__restrict__ cell_start, int *__restrict__ cell_atoms, int mbins, int nlocal,
                            const int *__restrict__ tag, const double4 *__restrict__ xq) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= mbins) return;
  int s = cell_start[b], e = cell_start[b + 1];
  for (int a = s + 1; a < e; a++) {
    int key = cell_atoms[a], k = a - 1;
    while (k >= s && cell_before<BYTAG, BYX>(key, cell_atoms[k], nlocal, tag, xq)) { cell_atoms[k + 1] = cell_atoms[k]; k--; }
    cell_atoms[k + 1] = key;
  }
}

// counting sort of atoms [0,n) into cells; leaves atom_bin, cell_start[0..mbins(+1)], cell_atoms.  nown: atoms below
// it are owned (in-cell order key).  dead != nullptr: flagged atoms go to an extra bin behind the last one.
static int bin_range(b200md_ctx *c, int n, int nown, const int *dead) {
  cudaStream_t s = c->stream;
  const int nb = c->mbins + (dead ? 1 : 0);
  CUDA_TRY(cudaMemsetAsync(c->cell_count, 0, sizeof(int) * (nb + 1), s));
  CUDA_TRY(cudaMemsetAsync(c->cell_fill, 0, sizeof(int) * (nb + 1), s));
  BinParams p = make_bin_params(c);
  if (n > 0) k_bin_count<<<(n + 255) / 256, 256, 0, s>>>(c->xq, n, p, c->atom_bin, c->cell_count, c->d_flags, dead, c->mbins);
  TRY(dev_exclusive_scan(c, c->cell_count, c->cell_start, nb + 1));
  if (n > 0) k_bin_scatter<<<(n + 255) / 256, 256, 0, s>>>(c->atom_bin, n, c->cell_start, c->cell_fill, c->cell_atoms);
  const int nbk = (c->mbins + 255) / 256;
  if (c->cell_by_x) {
    if (c->h26.on) k_cell_sort<true, true><<<nbk, 256, 0, s>>>(c->cell_start, c->cell_atoms, c->mbins, nown, c->tag, c->xq);
    else k_cell_sort<false, true><<<nbk, 256, 0, s>>>(c->cell_start, c->cell_atoms, c->mbins, nown, c->tag, c->xq);
  } else {
    if (c->h26.on) k_cell_sort<true, false><<<nbk, 256, 0, s>>>(c->cell_start, c->cell_atoms, c->mbins, nown, c->tag, c->xq);
    else k_cell_sort<false, false><<<nbk, 256, 0, s>>>(c->cell_start, c->cell_atoms, c->mbins, nown, c->tag, c->xq);
  }
  c->launches += 3;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

__global__ void k_permute_owned(int n, int nmax, const int *__restrict__ perm, const double4 *__restrict__ xq,
                                double4 *__restrict__ xq_o, const double *__restrict__ v, double *__restrict__ v_o,
                                const int *__restrict__ tag, int *__restrict__ tag_o,
                                const int *__restrict__ image, int *__restrict__ image_o) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int s = perm[i];
  xq_o[i] = xq[s];
  v_o[i] = v[s];
  v_o[i + nmax] = v[s + nmax];
  v_o[i + 2 * (size_t)nmax] = v[s + 2 * (size_t)nmax];
  tag_o[i] = tag[s];
  image_o[i] = image[s];
}

int neigh_sort_owned(b200md_ctx *c) {
  // after a migration of the direct halo c->nlocal still counts the atoms that left: they are flagged dead, sorted
  // behind the last bin and dropped here
  const bool drop = c->h26.on && c->h26.dead;
  const int ntot = c->nlocal;
  TRY(bin_range(c, ntot, ntot, drop ? c->h26.dead : nullptr));
  const int n = drop ? ntot - c->h26.ndead : ntot;
  if (drop) { c->nlocal = n; c->h26.ndead = 0; }
  if (n == 0) return B200MD_OK;
  k_permute_owned<<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->nmax, c->cell_atoms, c->xq, c->xq_tmp, c->v,
                                                         c->v_tmp, c->tag, c->tag_tmp, c->image, c->image_tmp);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  std::swap(c->xq, c->xq_tmp);
  std::swap(c->v, c->v_tmp);
  std::swap(c->tag, c->tag_tmp);
  std::swap(c->image, c->image_tmp);
  return B200MD_OK;
}

__global__ void k_gather_cell_xq(int n, const int *__restrict__ cell_atoms, const double4 *__restrict__ xq,
                                 double4 *__restrict__ cell_xq) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  cell_xq[i] = xq[cell_atoms[i]];
}

int neigh_bin_all(b200md_ctx *c) {
  const int n = c->nlocal + c->nghost;
  TRY(bin_range(c, n, c->nlocal, nullptr));
  if (n > 0) {
    k_gather_cell_xq<<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->cell_atoms, c->xq, c->cell_xq);
    c->launches++;
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

__global__ void k_save_xhold(int n, int nmax, const double4 *__restrict__ xq, double *__restrict__ xhold) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 q = xq[i];
  xhold[i] = q.x; xhold[i + nmax] = q.y; xhold[i + 2 * (size_t)nmax] = q.z;
}
int neigh_save_xhold(b200md_ctx *c) {
  if (c->nlocal == 0) return B200MD_OK;
  k_save_xhold<<<(c->nlocal + 255) / 256, 256, 0, c->stream>>>(c->nlocal, c->nmax, c->xq, c->xhold);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// neighbor list build
// ---------------------------------------------------------------------------------------
#ifndef BUILD_THREADS
#define BUILD_THREADS 128
#endif
#define MAX_ROWS 128
#define MAX_TYPES_SMEM 8

// special-bond handling of the molecular builders (NPairHalfBinNewton / NPairFullBin,
// src/npair_half_bin_newton.cpp:98-113, find_special src/npair.h:125-160)
struct MolParams {
  const int *tag, *spec_off, *spec_tag, *spec_which;
  int flag1, flag2, flag3;
  double hx, hy, hz;   // half box lengths; <= 0 for a non-periodic dimension
};

template <int TPA, bool HALF, bool MOL>
__global__ void __launch_bounds__(BUILD_THREADS)
k_build(int nlocal, const double4 *__restrict__ xq, const int *__restrict__ atom_bin,
        const int *__restrict__ cell_start, const int *__restrict__ cell_atoms,
        const double4 *__restrict__ cell_xq, int nrows, const int *__restrict__ row_off,
        const int *__restrict__ row_sx, const int *__restrict__ row_flag,
        const double *__restrict__ cutneighsq, int ntypes, int *__restrict__ neigh,
        int *__restrict__ numneigh, int pitch, int *flags, int *__restrict__ tile_ghost, int F, MolParams M) {
  // TPA lanes scan the candidates of one atom; the list is written in the layout of the force
  // kernels, which use F lanes per atom (AF atoms per warp tile)
  constexpr int A = 32 / TPA;
  const int AF = 32 / F;
  __shared__ int s_off[MAX_ROWS], s_sx[MAX_ROWS], s_flag[MAX_ROWS];
  __shared__ double s_cut[(MAX_TYPES_SMEM + 1) * (MAX_TYPES_SMEM + 1)];
  for (int r = threadIdx.x; r < nrows; r += blockDim.x) { s_off[r] = row_off[r]; s_sx[r] = row_sx[r]; s_flag[r] = row_flag[r]; }
  const int nt1 = ntypes + 1;
  const bool cut_smem = ntypes <= MAX_TYPES_SMEM;
  if (cut_smem) for (int r = threadIdx.x; r < nt1 * nt1; r += blockDim.x) s_cut[r] = cutneighsq[r];
  __syncthreads();

  const int lane = threadIdx.x & 31;
  const int sub = lane / TPA, l = lane % TPA;
  const unsigned submask_all = (TPA == 32) ? 0xffffffffu : (((1u << TPA) - 1u) << (sub * TPA));
  const unsigned below = submask_all & ((1u << lane) - 1u);
  const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int tile = warp_global;           // one warp = one tile of A atoms
  const int i = tile * A + sub;
  const bool active = i < nlocal;
  int *out = neigh + (size_t)(i / AF) * pitch * AF + (i % AF) * F;

  double4 qi = make_double4(0, 0, 0, 0);
  int ibin = 0, itype = 1;
  int s0 = 0, s1 = 0;
  if (active) { qi = xq[i]; ibin = atom_bin[i]; itype = xq_type(qi); }
  if (MOL && active) { const int ti = M.tag[i]; s0 = M.spec_off[ti]; s1 = M.spec_off[ti + 1]; }
  const double *cutrow = cut_smem ? &s_cut[itype * nt1] : &cutneighsq[itype * nt1];
  int cnt = 0;
  bool has_ghost = false;

  for (int r = 0; r < nrows; r++) {
    int lo = 0, hi = 0;
    const int role = s_flag[r];
    if (active) {
      const int b = ibin + s_off[r];
      if (HALF && role == 2) { lo = cell_start[b]; hi = cell_start[b + s_sx[r] + 1]; }   // own bin + dx>0
      else { lo = cell_start[b - s_sx[r]]; hi = cell_start[b + s_sx[r] + 1]; }
    }
    int own_end = 0;
    if (HALF && role == 2 && active) own_end = cell_start[ibin + 1];
    if (TPA == 1) {
      // one lane per atom: no votes, a private counter, and the next candidate in flight while this one is tested
      if (lo >= hi) continue;
      double4 qj = ld_xq(&cell_xq[lo]);
      int j = __ldg(&cell_atoms[lo]);
      for (int cidx = lo; cidx < hi; cidx++) {
        const int nx = min(cidx + 1, hi - 1);
        const double4 qn = ld_xq(&cell_xq[nx]);
        const int jn = __ldg(&cell_atoms[nx]);
        const double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
        bool keep = rsq_nofma(dx, dy, dz) <= cutrow[xq_type(qj)];
        if (keep) {
          if (HALF) {
            if (role == 2 && cidx < own_end) {
              if (j <= i) keep = false;
              else if (j >= nlocal) {
                if (qj.z < qi.z) keep = false;
                else if (qj.z == qi.z) {
                  if (qj.y < qi.y) keep = false;
                  else if (qj.y == qi.y && qj.x < qi.x) keep = false;
                }
              }
            }
          } else if (j == i) keep = false;
          if (MOL && keep && s1 > s0) {
            const int tj = M.tag[j];
            for (int sp = s0; sp < s1; sp++) {
              if (M.spec_tag[sp] == tj) {
                const int wh = M.spec_which[sp];
                const int fl = wh == 1 ? M.flag1 : (wh == 2 ? M.flag2 : M.flag3);
                if (fl == 0) {
                  const bool far = (M.hx > 0 && fabs(dx) > M.hx) || (M.hy > 0 && fabs(dy) > M.hy) || (M.hz > 0 && fabs(dz) > M.hz);
                  if (!far) keep = false;
                }
                break;
              }
            }
          }
          if (keep) {
            has_ghost |= (j >= nlocal);
            if (cnt < pitch) out[(cnt / F) * 32 + (cnt % F)] = j;
            cnt++;
          }
        }
        qj = qn; j = jn;
      }
      continue;
    }
    int cidx = lo + l;
    while (__any_sync(0xffffffffu, cidx < hi)) {
      bool keep = false;
      int j = -1;
      if (cidx < hi) {
        const double4 qj = ld_xq(&cell_xq[cidx]);
        j = __ldg(&cell_atoms[cidx]);
        const double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
        const double rsq = rsq_nofma(dx, dy, dz);
        keep = rsq <= cutrow[xq_type(qj)];
        if (HALF) {
          if (role == 2 && cidx < own_end) {
            // rest of i's own bin (npair_half_bin_atomonly_newton.cpp:61-69)
            if (j <= i) keep = false;
            else if (j >= nlocal) {
              if (qj.z < qi.z) keep = false;
              else if (qj.z == qi.z) {
                if (qj.y < qi.y) keep = false;
                else if (qj.y == qi.y && qj.x < qi.x) keep = false;
              }
            }
          }
        } else {
          if (j == i) keep = false;
        }
        if (MOL && keep && s1 > s0) {
          // find_special: excluded (flag 0) unless the pair is more than half a box apart
          const int tj = M.tag[j];
          for (int sp = s0; sp < s1; sp++) {
            if (M.spec_tag[sp] == tj) {
              const int wh = M.spec_which[sp];
              const int fl = wh == 1 ? M.flag1 : (wh == 2 ? M.flag2 : M.flag3);
              if (fl == 0) {
                const bool far = (M.hx > 0 && fabs(dx) > M.hx) || (M.hy > 0 && fabs(dy) > M.hy) || (M.hz > 0 && fabs(dz) > M.hz);
                if (!far) keep = false;
              }
              break;
            }
          }
        }
      }
      const unsigned m = __ballot_sync(0xffffffffu, keep) & submask_all;
      if (keep) {
        has_ghost |= (j >= nlocal);
        const int k = cnt + __popc(m & below);
        if (k < pitch) out[(k / F) * 32 + (k % F)] = j;
      }
      cnt += __popc(m);
      cidx += TPA;
    }
  }
  if (active && l == 0) {
    numneigh[i] = cnt;
    if (cnt > pitch) atomicMax(&flags[1], cnt);
  }
  // tiles whose atoms have no ghost neighbor can be evaluated while the halo is in flight
  // (tile_ghost is cleared before the launch; all writers store the same value)
  if (tile_ghost && has_ghost) tile_ghost[i / AF] = 1;
}


// ---------------------------------------------------------------------------------------
// cluster-pair list (atomic systems): CL = 4 consecutive owned atoms of the bin-sorted order form an
// i-cluster with ONE j-list; entry = j | mask << 28, bit k of the mask set iff (i0+k, j) is a pair of the
// reference's exact per-atom list (rsq <= cutneighsq[itype][jtype] and the full / half-newton membership
// rule).  The force kernels gather each j once per cluster, evaluate up to four pairs from registers and
// issue one RED triple per (cluster, j) instead of one per pair; the exact per-atom lists are recovered
// from the masks (b200md_neigh_export).  The union of the four stencils is scanned once: the atoms are
// bin-sorted, so their stencil rows overlap and (same bin row) merge into one contiguous range of
// cell_xq.  Membership in atom k's stencil needs no test of its own: a candidate outside it is farther
// than cutneighmax (NStencil keeps every bin closer than that, nstencil_full_bin_3d.cpp:35).
// Half/newton rule (npair_half_bin_atomonly_newton.cpp:55-104) in the device's order — owned atoms sorted
// by bin, ascending inside a bin, ghosts after the owned atoms of a cell:
//   owned j : j in a later bin, or in i's bin behind i            <=>  j > i
//   ghost j : j in a later bin (cell position >= end of i's bin), or in i's bin and not below i in
//             (z, y, x) order
// ---------------------------------------------------------------------------------------
template <bool HALF, bool ONETYPE>
__global__ void __launch_bounds__(BUILD_THREADS)
k_build_cluster(int nlocal, int ncl, const double4 *__restrict__ xq, const int *__restrict__ atom_bin,
                const int *__restrict__ cell_start, const int *__restrict__ cell_atoms,
                const double4 *__restrict__ cell_xq, int nrows, const int *__restrict__ row_off,
                const int *__restrict__ row_sx, const int *__restrict__ row_flag,
                const double *__restrict__ cutneighsq, int ntypes, int mbinx, int *__restrict__ neigh,
                int *__restrict__ numneigh, int pitch, int *flags, int *__restrict__ tile_ghost, int L) {
  __shared__ int s_off[MAX_ROWS], s_sx[MAX_ROWS], s_flag[MAX_ROWS];
  __shared__ double s_cut[(MAX_TYPES_SMEM + 1) * (MAX_TYPES_SMEM + 1)];
  for (int r = threadIdx.x; r < nrows; r += blockDim.x) { s_off[r] = row_off[r]; s_sx[r] = row_sx[r]; s_flag[r] = row_flag[r]; }
  const int nt1 = ntypes + 1;
  const bool cut_smem = ntypes <= MAX_TYPES_SMEM;
  if (cut_smem) for (int r = threadIdx.x; r < nt1 * nt1; r += blockDim.x) s_cut[r] = cutneighsq[r];
  __syncthreads();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= ncl) return;
  const int C = 32 / L;
  int *out = neigh + (size_t)(c / C) * pitch * C + (c % C) * L;
  const int i0 = c * CL;
  const int ni = min(CL, nlocal - i0);
  double xi[CL], yi[CL], zi[CL];
  int bi[CL], own_lo[CL], own_hi[CL];
  const double *cutrow[CL];
  const double cut1 = ONETYPE ? cutneighsq[nt1 + 1] : 0.0;
#pragma unroll
  for (int k = 0; k < CL; k++) {
    const int i = i0 + min(k, ni - 1);
    const double4 q = xq[i];
    // an absent atom of the last cluster sits far away: no candidate is ever within its cutoff
    xi[k] = k < ni ? q.x : 1e150; yi[k] = q.y; zi[k] = q.z;
    bi[k] = atom_bin[i];
    own_lo[k] = cell_start[bi[k]]; own_hi[k] = cell_start[bi[k] + 1];
    const int it = xq_type(q);
    cutrow[k] = cut_smem ? &s_cut[it * nt1] : &cutneighsq[it * nt1];
  }
  int cnt = 0;
  bool has_ghost = false;
  // The atoms of a cluster are consecutive in bin order, so they normally sit in one row of bins (same iy, iz) and
  // the union of their stencil rows is one contiguous range per (dy,dz).  A cluster that straddles the end of a bin
  // row is handled as several groups, one per bin row, each scanning for its own atoms only: a j reached from two
  // groups then appears in two entries with disjoint masks, never as a duplicate pair.
  for (int g0 = 0; g0 < CL;) {
    int g1 = g0 + 1;
    while (g1 < CL && bi[g1] / mbinx == bi[g0] / mbinx) g1++;    // bins ascend with k
    const unsigned gmask = ((1u << g1) - 1u) & ~((1u << g0) - 1u);
    const int bmin = bi[g0], bmax = bi[g1 - 1];
    for (int r = 0; r < nrows; r++) {
      const int off = s_off[r], sx = s_sx[r];
      const int back = (HALF && s_flag[r] == 2) ? 0 : sx;        // own row of the half stencil: own bin and dx > 0
      const int lo = cell_start[bmin + off - back], hi = cell_start[bmax + off + sx + 1];
      if (lo >= hi) continue;
      // the thread walks its candidates one by one; the next record is in flight while this one is tested
      double4 qj = ld_xq(&cell_xq[lo]);
      int j = __ldg(&cell_atoms[lo]);
      for (int cidx = lo; cidx < hi; cidx++) {
        const int nx = min(cidx + 1, hi - 1);
        const double4 qn = ld_xq(&cell_xq[nx]);
        const int jn = __ldg(&cell_atoms[nx]);
        const int jt = ONETYPE ? 1 : xq_type(qj);
        unsigned m = 0;
#pragma unroll
        for (int a = 0; a < CL; a++) {
          const double dx = xi[a] - qj.x, dy = yi[a] - qj.y, dz = zi[a] - qj.z;
          m |= (rsq_nofma(dx, dy, dz) <= (ONETYPE ? cut1 : cutrow[a][jt]) ? 1u : 0u) << a;
        }
        m &= gmask;
        if (m) {
          const int d = j - i0;
          if (HALF) {
            if (j < nlocal) m &= d <= 0 ? 0u : (d >= CL ? CL_FULL : (1u << d) - 1u);   // owned j pairs with the atoms before it
            else {
#pragma unroll
              for (int a = 0; a < CL; a++) {
                bool keep = true;
                if (cidx < own_hi[a]) {          // not in a later bin than atom a's
                  if (cidx < own_lo[a]) keep = false;
                  else if (qj.z < zi[a]) keep = false;
                  else if (qj.z == zi[a]) {
                    if (qj.y < yi[a]) keep = false;
                    else if (qj.y == yi[a] && qj.x < xi[a]) keep = false;
                  }
                }
                if (!keep) m &= ~(1u << a);
              }
            }
          } else if (d >= 0 && d < CL) m &= ~(1u << d);
          if (m) {
            if (cnt < pitch) out[(cnt / L) * 32 + (cnt % L)] = (int)((unsigned)j | (m << CL_SHIFT));
            cnt++;
            has_ghost |= (j >= nlocal);
          }
        }
        qj = qn; j = jn;
      }
    }
    g0 = g1;
  }
  numneigh[c] = cnt;
  if (cnt > pitch) atomicMax(&flags[1], cnt);
  if (tile_ghost && has_ghost) tile_ghost[c / C] = 1;
}

// Small systems: the same cluster build with T lanes per cluster (one thread per cluster leaves most SMs empty below
// ~10^5 atoms and walks ~200 candidates serially).  The T lanes stride over each candidate range; kept entries are
// compacted with a ballot over the group's lanes, so the entry order (ascending candidates) is the same as with T = 1.
template <bool HALF, bool ONETYPE, int T>
__global__ void __launch_bounds__(BUILD_THREADS)
k_build_cluster_mt(int nlocal, int ncl, const double4 *__restrict__ xq, const int *__restrict__ atom_bin,
                   const int *__restrict__ cell_start, const int *__restrict__ cell_atoms,
                   const double4 *__restrict__ cell_xq, int nrows, const int *__restrict__ row_off,
                   const int *__restrict__ row_sx, const int *__restrict__ row_flag,
                   const double *__restrict__ cutneighsq, int ntypes, int mbinx, int *__restrict__ neigh,
                   int *__restrict__ numneigh, int pitch, int *flags, int *__restrict__ tile_ghost, int L) {
  __shared__ int s_off[MAX_ROWS], s_sx[MAX_ROWS], s_flag[MAX_ROWS];
  __shared__ double s_cut[(MAX_TYPES_SMEM + 1) * (MAX_TYPES_SMEM + 1)];
  for (int r = threadIdx.x; r < nrows; r += blockDim.x) { s_off[r] = row_off[r]; s_sx[r] = row_sx[r]; s_flag[r] = row_flag[r]; }
  const int nt1 = ntypes + 1;
  const bool cut_smem = ntypes <= MAX_TYPES_SMEM;
  if (cut_smem) for (int r = threadIdx.x; r < nt1 * nt1; r += blockDim.x) s_cut[r] = cutneighsq[r];
  __syncthreads();
  const int gt = blockIdx.x * blockDim.x + threadIdx.x;
  const int c = gt / T, l = gt % T;
  if (c >= ncl) return;                     // whole groups leave together
  const int lane = threadIdx.x & 31;
  const unsigned gmask_lanes = (T == 32) ? 0xffffffffu : (((1u << T) - 1u) << (lane - l));
  const unsigned below = gmask_lanes & ((1u << lane) - 1u);
  const int C = 32 / L;
  int *out = neigh + (size_t)(c / C) * pitch * C + (c % C) * L;
  const int i0 = c * CL;
  const int ni = min(CL, nlocal - i0);
  double xi[CL], yi[CL], zi[CL];
  int bi[CL], own_lo[CL], own_hi[CL];
  const double *cutrow[CL];
  const double cut1 = ONETYPE ? cutneighsq[nt1 + 1] : 0.0;
#pragma unroll
  for (int k = 0; k < CL; k++) {
    const int i = i0 + min(k, ni - 1);
    const double4 q = xq[i];
    xi[k] = k < ni ? q.x : 1e150; yi[k] = q.y; zi[k] = q.z;
    bi[k] = atom_bin[i];
    own_lo[k] = cell_start[bi[k]]; own_hi[k] = cell_start[bi[k] + 1];
    const int it = xq_type(q);
    cutrow[k] = cut_smem ? &s_cut[it * nt1] : &cutneighsq[it * nt1];
  }
  int cnt = 0;
  bool has_ghost = false;
  for (int g0 = 0; g0 < CL;) {
    int g1 = g0 + 1;
    while (g1 < CL && bi[g1] / mbinx == bi[g0] / mbinx) g1++;
    const unsigned gmask = ((1u << g1) - 1u) & ~((1u << g0) - 1u);
    const int bmin = bi[g0], bmax = bi[g1 - 1];
    for (int r = 0; r < nrows; r++) {
      const int off = s_off[r], sx = s_sx[r];
      const int back = (HALF && s_flag[r] == 2) ? 0 : sx;
      const int lo = cell_start[bmin + off - back], hi = cell_start[bmax + off + sx + 1];
      for (int base = lo; base < hi; base += T) {
        const int cidx = base + l;
        unsigned m = 0;
        int j = 0;
        if (cidx < hi) {
          const double4 qj = ld_xq(&cell_xq[cidx]);
          j = __ldg(&cell_atoms[cidx]);
          const int jt = ONETYPE ? 1 : xq_type(qj);
#pragma unroll
          for (int a = 0; a < CL; a++) {
            const double dx = xi[a] - qj.x, dy = yi[a] - qj.y, dz = zi[a] - qj.z;
            m |= (rsq_nofma(dx, dy, dz) <= (ONETYPE ? cut1 : cutrow[a][jt]) ? 1u : 0u) << a;
          }
          m &= gmask;
          if (m) {
            const int d = j - i0;
            if (HALF) {
              if (j < nlocal) m &= d <= 0 ? 0u : (d >= CL ? CL_FULL : (1u << d) - 1u);
              else {
#pragma unroll
                for (int a = 0; a < CL; a++) {
                  bool keep = true;
                  if (cidx < own_hi[a]) {
                    if (cidx < own_lo[a]) keep = false;
                    else if (qj.z < zi[a]) keep = false;
                    else if (qj.z == zi[a]) {
                      if (qj.y < yi[a]) keep = false;
                      else if (qj.y == yi[a] && qj.x < xi[a]) keep = false;
                    }
                  }
                  if (!keep) m &= ~(1u << a);
                }
              }
            } else if (d >= 0 && d < CL) m &= ~(1u << d);
          }
        }
        const unsigned kept = __ballot_sync(gmask_lanes, m != 0) & gmask_lanes;
        if (m) {
          const int k = cnt + __popc(kept & below);
          if (k < pitch) out[(k / L) * 32 + (k % L)] = (int)((unsigned)j | (m << CL_SHIFT));
          has_ghost |= (j >= nlocal);
        }
        cnt += __popc(kept);
      }
    }
    g0 = g1;
  }
  has_ghost = __any_sync(gmask_lanes, has_ghost);
  if (l == 0) {
    numneigh[c] = cnt;
    if (cnt > pitch) atomicMax(&flags[1], cnt);
    if (tile_ghost && has_ghost) tile_ghost[c / C] = 1;
  }
}

template <int TPA, bool HALF, bool MOL>
static void launch_build3(b200md_ctx *c, int nblocks, const MolParams &M) {
  k_build<TPA, HALF, MOL><<<nblocks, BUILD_THREADS, 0, c->stream>>>(
      c->nlocal, c->xq, c->atom_bin, c->cell_start, c->cell_atoms, c->cell_xq, c->nrows, c->row_off,
      c->row_sx, c->row_flag, c->cutneighsq, c->ntypes, c->neigh, c->numneigh, c->pitch, c->d_flags, c->tile_ghost, c->tpa, M);
}
template <int TPA>
static void launch_build(b200md_ctx *c) {
  constexpr int A = 32 / TPA;
  const int ntiles = (c->nlocal + A - 1) / A;
  const int warps_per_block = BUILD_THREADS / 32;
  const int nblocks = (ntiles + warps_per_block - 1) / warps_per_block;
  if (nblocks == 0) return;
  const Topology &t = c->topo;
  MolParams M = {c->tag, t.spec_off, t.spec_tag, t.spec_which, t.special_flag[1], t.special_flag[2], t.special_flag[3],
                 c->periodic[0] ? 0.5 * c->prd[0] : -1.0, c->periodic[1] ? 0.5 * c->prd[1] : -1.0,
                 c->periodic[2] ? 0.5 * c->prd[2] : -1.0};
  const bool half = c->list_kind == B200MD_LIST_HALF_NEWTON;
  if (t.molecular) { if (half) launch_build3<TPA, true, true>(c, nblocks, M); else launch_build3<TPA, false, true>(c, nblocks, M); }
  else { if (half) launch_build3<TPA, true, false>(c, nblocks, M); else launch_build3<TPA, false, false>(c, nblocks, M); }
}

static int ensure_neigh_storage(b200md_ctx *c) {
  const int A = 32 / c->tpa;
  const size_t ntiles = (size_t)(list_units(c) + A - 1) / A + 1;
  const size_t need = ntiles * (size_t)c->pitch * A;
  if (need > c->neigh_cap) {
    if (c->neigh) CUDA_TRY(cudaFree(c->neigh));
    c->neigh_cap = need + need / 8;
    CUDA_TRY(cudaMalloc(&c->neigh, c->neigh_cap * sizeof(int)));
  }
  if ((int)ntiles + 1 > c->tile_cap) {
    int *p[] = {c->tile_ghost, c->tile_pos, c->tiles_int, c->tiles_bnd};
    for (int *q : p) if (q) CUDA_TRY(cudaFree(q));
    c->tile_cap = (int)ntiles + 1 + (int)ntiles / 8;
    CUDA_TRY(cudaMalloc(&c->tile_ghost, sizeof(int) * c->tile_cap));
    CUDA_TRY(cudaMalloc(&c->tile_pos, sizeof(int) * c->tile_cap));
    CUDA_TRY(cudaMalloc(&c->tiles_int, sizeof(int) * c->tile_cap));
    CUDA_TRY(cudaMalloc(&c->tiles_bnd, sizeof(int) * c->tile_cap));
  }
  return B200MD_OK;
}

__global__ void k_split_tiles(int ntiles, const int *__restrict__ flag, const int *__restrict__ pos,
                              int *__restrict__ tiles_int, int *__restrict__ tiles_bnd) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= ntiles) return;
  if (flag[t]) tiles_bnd[pos[t]] = t; else tiles_int[t - pos[t]] = t;
}

// interior / boundary tile lists for compute-halo overlap (multi-rank runs only)
static int neigh_split_tiles(b200md_ctx *c) {
  c->n_tiles_int = c->n_tiles_bnd = 0;
  if (c->all_self || c->nlocal == 0) return B200MD_OK;
  const int ntiles = list_tiles(c);
  CUDA_TRY(cudaMemsetAsync(c->tile_ghost + ntiles, 0, sizeof(int), c->stream));
  TRY(dev_exclusive_scan(c, c->tile_ghost, c->tile_pos, ntiles + 1));
  k_split_tiles<<<(ntiles + 255) / 256, 256, 0, c->stream>>>(ntiles, c->tile_ghost, c->tile_pos, c->tiles_int, c->tiles_bnd);
  c->launches++;
  CUDA_TRY(cudaMemcpyAsync(c->h_flags + 5, c->tile_pos + ntiles, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  c->n_tiles_bnd = c->h_flags[5];
  c->n_tiles_int = ntiles - c->n_tiles_bnd;
  return B200MD_OK;
}

int neigh_build_list(b200md_ctx *c) {
  if (c->nrows > MAX_ROWS) { b200md_set_error("b200md: stencil has too many rows"); return B200MD_ERR; }
  // atomic systems use the cluster-pair list; molecular ones (special-bond exclusions, short lists) the per-atom list
  const bool cluster = !c->topo.molecular && (c->use_cluster < 0 ? c->pair_style == B200MD_PAIR_LJ_CUT : c->use_cluster != 0);
  if (cluster != c->cluster) { c->cluster = cluster; c->pitch = 0; }
  // lanes per list unit, measured on B200 (profiles/): lj/cut 4 per cluster, 2 per atom on the per-atom list; eam 1 per atom
  // small systems (fewer lanes than the SMs can hold) are latency-bound: more lanes per unit shorten the serial chains
  const bool small = c->nlocal < 300000;
  if (!c->tpa_user) c->tpa = cluster ? (small ? 8 : 4) : (c->pair_style == B200MD_PAIR_EAM ? (small ? 4 : 1) : 2);
  if (!c->tpa_build_user) c->tpa_build = (small && c->pair_style == B200MD_PAIR_EAM) ? 8 : 1;
  if (cluster && c->nlocal + c->nghost > CL_JMASK) { b200md_set_error("b200md: more than 2^28 atoms on one rank"); return B200MD_ERR; }
  if (c->pitch == 0) {
    // first guess from the number density: full sphere of cutneighmax, +35 % slack; a cluster's j-list is the union
    // of four such spheres around atoms about one bin apart (measured 1.7-1.9 x one sphere)
    double vol = (c->subhi[0] - c->sublo[0]) * (c->subhi[1] - c->sublo[1]) * (c->subhi[2] - c->sublo[2]);
    double dens = vol > 0 ? c->nlocal / vol : 1.0;
    double est = 4.18879 * c->cutneighmax * c->cutneighmax * c->cutneighmax * dens;
    if (c->list_kind == B200MD_LIST_HALF_NEWTON) est *= 0.5;
    if (cluster) est *= 2.0;
    int p = (int)(est * 1.35) + 16;
    c->pitch = (p + 31) / 32 * 32;
  }
  for (int attempt = 0; attempt < 8; attempt++) {
    TRY(ensure_neigh_storage(c));
    c->h_flags[1] = 0;
    CUDA_TRY(cudaMemsetAsync(c->d_flags + 1, 0, sizeof(int), c->stream));
    CUDA_TRY(cudaMemsetAsync(c->tile_ghost, 0, sizeof(int) * (list_tiles(c) + 1), c->stream));
    if (cluster) {
      const int ncl = list_units(c);
      if (ncl) {
        const int nb = (ncl + BUILD_THREADS - 1) / BUILD_THREADS;
        const bool half = c->list_kind == B200MD_LIST_HALF_NEWTON, one = c->ntypes == 1;
#define B200_LAUNCH_BUILD_CL(H, O)                                                                                             \
  k_build_cluster<H, O><<<nb, BUILD_THREADS, 0, c->stream>>>(c->nlocal, ncl, c->xq, c->atom_bin, c->cell_start, c->cell_atoms, \
      c->cell_xq, c->nrows, c->row_off, c->row_sx, c->row_flag, c->cutneighsq, c->ntypes, c->bin.mbinx, c->neigh, c->numneigh,  \
      c->pitch, c->d_flags, c->tile_ghost, c->tpa)
#define B200_LAUNCH_BUILD_CL_MT(H, O)                                                                                          \
  k_build_cluster_mt<H, O, 8><<<(ncl * 8 + BUILD_THREADS - 1) / BUILD_THREADS, BUILD_THREADS, 0, c->stream>>>(                    \
      c->nlocal, ncl, c->xq, c->atom_bin, c->cell_start, c->cell_atoms, c->cell_xq, c->nrows, c->row_off, c->row_sx, c->row_flag, \
      c->cutneighsq, c->ntypes, c->bin.mbinx, c->neigh, c->numneigh, c->pitch, c->d_flags, c->tile_ghost, c->tpa)
        const bool mt = c->tpa_build_user ? c->tpa_build > 1 : small;   // 8 lanes per cluster below ~3 10^5 atoms
        if (mt) {
          if (half) { if (one) B200_LAUNCH_BUILD_CL_MT(true, true); else B200_LAUNCH_BUILD_CL_MT(true, false); }
          else { if (one) B200_LAUNCH_BUILD_CL_MT(false, true); else B200_LAUNCH_BUILD_CL_MT(false, false); }
        } else if (half) { if (one) B200_LAUNCH_BUILD_CL(true, true); else B200_LAUNCH_BUILD_CL(true, false); }
        else { if (one) B200_LAUNCH_BUILD_CL(false, true); else B200_LAUNCH_BUILD_CL(false, false); }
#undef B200_LAUNCH_BUILD_CL
#undef B200_LAUNCH_BUILD_CL_MT
      }
    } else
    switch (c->tpa_build) {
      case 1: launch_build<1>(c); break;
      case 2: launch_build<2>(c); break;
      case 4: launch_build<4>(c); break;
      case 8: launch_build<8>(c); break;
      case 16: launch_build<16>(c); break;
      case 32: launch_build<32>(c); break;
      default: b200md_set_error("b200md: build lanes per atom must be 1, 2, 4, 8, 16 or 32"); return B200MD_ERR;
    }
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(c->h_flags, c->d_flags, 5 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    if (c->h_flags[2]) { b200md_set_error("Non-numeric positions - simulation unstable"); return B200MD_ERR; }
    // bond kernel flags of the steps since the last rebuild (bond.cu): the reference stops at the rebuild
    // (NTopoBondAll::build, src/ntopo_bond_all.cpp:37-80) or at the step itself (BondFENE::compute)
    if (c->h_flags[4] & 1) { b200md_set_error("Bond atoms missing"); return B200MD_ERR; }
    if (c->h_flags[4] & 4) { b200md_set_error("Bad FENE bond"); return B200MD_ERR; }
    if (c->h_flags[4]) CUDA_TRY(cudaMemsetAsync(c->d_flags + 4, 0, sizeof(int), c->stream));   // warning bit: reported once
    if (c->h_flags[1] == 0) return neigh_split_tiles(c);
    // overflow: regrow the pitch like the reference's Kokkos path (npair_kokkos.cpp:302-316)
    int p = (int)(c->h_flags[1] * 1.2) + 8;
    c->pitch = (p + 31) / 32 * 32;
  }
  b200md_set_error("Neighbor list overflow, boost neigh_modify one");
  return B200MD_ERR;
}
