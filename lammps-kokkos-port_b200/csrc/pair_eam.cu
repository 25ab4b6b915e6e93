// pair_eam.cu — pair_style eam / eam/alloy force evaluation (same compute for both; the
// file formats differ only in how the spline tables are produced on the host side).
//
// Reference: PairEAM::compute (/root/reference/src/MANYBODY/pair_eam.cpp:142-328)
//   pass 1 (:187-220)  rho_i += rhor_spline[type2rhor[jtype][itype]][m] value cubic at p
//   embed  (:231-247)  fp_i = F'(rho_i); phi = F(rho_i) (+ linear extrapolation past rhomax)
//   comm   (:251)      forward_comm(fp)
//   pass 3 (:257-325)  psip = fp_i*rhojp + fp_j*rhoip + phip; fpair = -scale*psip/r
// Half list / newton on (the reference's CPU default) is also provided:
//   k_eam_rho_half (rho_i and rho_j per listed pair, FP64 RED) -> reverse_comm(rho)
//   (pair_eam.cpp:224) -> k_eam_embed -> forward_comm(fp) -> k_eam_force<...,HALF> (f_j via RED)
// Default B200 design: full list / newton off, so pass 1 needs no reverse_comm of rho and pass 1 and
// the embedding are one kernel (what the reference's Kokkos port calls KernelAB,
// src/KOKKOS/pair_eam_kokkos.cpp:839-921): TPA lanes gather, shuffle-reduce rho_i, lane 0
// evaluates F and F'.  Tables are repacked so that each lookup is whole 32-byte sectors:
//   rhoA[t][m] = {c3,c4,c5,c6}   density value cubic      (1 sector)
//   z2rV[t][m] = {c3,c4,c5,c6}   pair-term value cubic    (1 sector)
//   tabC[m]    = {rhoA row, z2rV row}  when there is one element (64 contiguous bytes)
// The derivative rows c0..c2 of the reference's 7-wide splines are not stored: interpolate()
// defines them as c0=3*c3/delta, c1=2*c4/delta, c2=c5/delta (pair_eam.cpp:792-796), so the
// kernels evaluate ((3*c3*p + 2*c4)*p + c5)*rdr — same polynomial, rounding differs at 1e-16.
// This cuts the per-pair table traffic of pass 3 from 96 to 64 bytes; the L1/shared data
// pipe is what bounds these kernels (ncu: l1tex throughput 80-88 %).
// One-element runs also gather a packed {x,y,z,fp} record (xfp, one sector) in pass 3 instead
// of {x,y,z,type} + fp[j] (two sectors).  tabC is staged into shared memory when it fits
// (Cu_u3: nr=499 -> 32 KB), otherwise read through L1/L2 (Cu_mishin1: nr=10001 -> 640 KB).
#include "ctx.h"
#include "util.cuh"

#define EAM_THREADS 256
// tuning macros (see pair_lj.cu): unroll depth and minimum resident blocks per SM of the pair loops
// measured best on B200: density pass 32 registers / 8 blocks per SM / no unroll; force pass 64
// registers / unroll 2 (it needs the registers: 40-register builds spill and lose 15 %)
#ifndef EAM_RHO_UNROLL
#define EAM_RHO_UNROLL 1
#endif
#ifndef EAM_RHO_MINBLOCKS
#define EAM_RHO_MINBLOCKS 8
#endif
// block size of the half-list density pass (it shares nothing between warps unless its table is staged): 2048 threads per
// SM as blocks of EAM_RHO_THREADS
#ifndef EAM_RHO_THREADS
#define EAM_RHO_THREADS 256
#endif
#define EAM_RHO_HALF_MINBLOCKS ((EAM_RHO_MINBLOCKS * 256 / EAM_RHO_THREADS) > 32 ? 32 : (EAM_RHO_MINBLOCKS * 256 / EAM_RHO_THREADS))
#ifndef EAM_FORCE_UNROLL
#define EAM_FORCE_UNROLL 2
#endif
#ifndef EAM_FORCE_MINBLOCKS
#define EAM_FORCE_MINBLOCKS 4
#endif

void pair_eam_free(b200md_ctx *c) {
  EamTables &e = c->eam;
  void *p[] = {e.frho, e.rhoA, e.z2rV, e.tabC, e.type2frho, e.type2rhor, e.type2z2r, e.scale, e.xfp, e.pmask};
  for (void *q : p) if (q) cudaFree(q);
  e = EamTables();
}

template <typename T>
static int upload(T *&dst, const std::vector<T> &h) {
  CUDA_TRY(cudaMalloc(&dst, sizeof(T) * h.size()));
  CUDA_TRY(cudaMemcpy(dst, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice));
  return B200MD_OK;
}

int pair_eam_config_impl(b200md_ctx *c, int ntypes, int nrho, int nr, double rdr, double rdrho, double rhomax,
                         double cutforcesq, int nfrho, const double *frho_spline, int nrhor,
                         const double *rhor_spline, int nz2r, const double *z2r_spline, const int *type2frho,
                         const int *type2rhor, const int *type2z2r, const double *scale) {
  pair_eam_free(c);
  EamTables &e = c->eam;
  e.ntypes = ntypes; e.nrho = nrho; e.nr = nr; e.nfrho = nfrho; e.nrhor = nrhor; e.nz2r = nz2r;
  e.rdr = rdr; e.rdrho = rdrho; e.rhomax = rhomax; e.cutforcesq = cutforcesq;
  const size_t rows = (size_t)nr + 1;
  std::vector<double> frho((size_t)nfrho * (nrho + 1) * 8, 0.0), rhoA((size_t)nrhor * rows * 4, 0.0),
      z2rV((size_t)nz2r * rows * 4, 0.0);
  for (int t = 0; t < nfrho; t++)
    for (int m = 0; m <= nrho; m++)
      for (int k = 0; k < 7; k++) frho[((size_t)t * (nrho + 1) + m) * 8 + k] = frho_spline[((size_t)t * (nrho + 1) + m) * 7 + k];
  for (int t = 0; t < nrhor; t++)
    for (size_t m = 0; m < rows; m++)
      for (int k = 0; k < 4; k++) rhoA[(t * rows + m) * 4 + k] = rhor_spline[(t * rows + m) * 7 + 3 + k];
  for (int t = 0; t < nz2r; t++)
    for (size_t m = 0; m < rows; m++)
      for (int k = 0; k < 4; k++) z2rV[(t * rows + m) * 4 + k] = z2r_spline[(t * rows + m) * 7 + 3 + k];
  const size_t nt = (size_t)(ntypes + 1) * (ntypes + 1);
  std::vector<int> t2f(type2frho, type2frho + ntypes + 1), t2r(type2rhor, type2rhor + nt), t2z(type2z2r, type2z2r + nt);
  std::vector<double> sc(nt, 1.0);
  if (scale) sc.assign(scale, scale + nt);
  e.one = (ntypes == 1);
  if (e.one) {
    const int tr = t2r[1 * 2 + 1], tz = t2z[1 * 2 + 1];
    std::vector<double> tabC(rows * 8);
    for (size_t m = 0; m < rows; m++)
      for (int k = 0; k < 4; k++) {
        tabC[m * 8 + k] = rhoA[(tr * rows + m) * 4 + k];
        tabC[m * 8 + 4 + k] = z2rV[(tz * rows + m) * 4 + k];
      }
    TRY(upload(e.tabC, tabC));
    e.scale11 = sc[1 * 2 + 1];
  }
  TRY(upload(e.frho, frho)); TRY(upload(e.rhoA, rhoA)); TRY(upload(e.z2rV, z2rV));
  TRY(upload(e.type2frho, t2f)); TRY(upload(e.type2rhor, t2r)); TRY(upload(e.type2z2r, t2z)); TRY(upload(e.scale, sc));
  return B200MD_OK;
}

constexpr int kRhoUnroll = EAM_RHO_UNROLL, kForceUnroll = EAM_FORCE_UNROLL;

__device__ __forceinline__ double4 ld4(const double *p) {
#if defined(B200MD_LD256)
  return ld_xq(reinterpret_cast<const double4 *>(p));   // one 256-bit load per 32-byte table row
#endif
  const double2 *q = reinterpret_cast<const double2 *>(p);
  const double2 a = q[0], b = q[1];
  return make_double4(a.x, a.y, b.x, b.y);
}

// ---- pass 1 + embedding -------------------------------------------------------------------
// ONE: single element (tabC, optional shared-memory staging, writes the packed xfp record)
template <int TPA, bool EV, bool ONE, bool SMEM>
__global__ void __launch_bounds__(EAM_THREADS, (EV || !ONE) ? 4 : EAM_RHO_MINBLOCKS)
k_eam_rho_fp(int nlocal, const double4 *__restrict__ xq, const int *__restrict__ neigh, const int *__restrict__ numneigh,
             int pitch, EamParams P, double *__restrict__ rho, double *__restrict__ fp, double4 *__restrict__ xfp,
             double *__restrict__ partials, const int *__restrict__ tile_list, int ntile_list) {
  constexpr int A = 32 / TPA;
  extern __shared__ double s_dyn[];
  const double *tab = ONE ? P.tabC : P.rhoA;
  if (ONE && SMEM) {
    const int n8 = (P.nr + 1) * 8;
    for (int k = threadIdx.x; k < n8; k += blockDim.x) s_dyn[k] = P.tabC[k];
    __syncthreads();
    tab = s_dyn;
  }
  const int lane = threadIdx.x & 31;
  const int sub = lane / TPA, l = lane % TPA;
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const bool in_list = !tile_list || w < ntile_list;
  const int tile = tile_list ? (in_list ? tile_list[w] : 0) : w;
  const int i = tile * A + sub;
  const bool active = in_list && i < nlocal;
  const int nt1 = P.ntypes + 1;
  double r_acc = 0;
  int itype = 1;
  double4 qi = make_double4(0, 0, 0, 0);
  if (active) {
    qi = xq[i];
    itype = ONE ? 1 : xq_type(qi);
    const int n = numneigh[i];
    const int *nb = neigh + (size_t)tile * pitch * A + sub * TPA + l;
    const int niter = (n + TPA - 1) / TPA;
#pragma unroll kRhoUnroll
    for (int it = 0; it < niter; it++) {
      if (it * TPA + l < n) {
        const int j = __ldg(nb + it * 32);
        const double4 qj = ld_xq(&xq[j]);
        const double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
        const double rsq = rsq_nofma(dx, dy, dz);
        if (rsq < P.cutforcesq) {
          double p = sqrt(rsq) * P.rdr + 1.0;
          int m = (int)p;
          m = min(m, P.nr - 1);
          p -= m;
          p = fmin(p, 1.0);
          double4 cf;
          if (ONE) cf = ld4(tab + (size_t)m * 8);
          else cf = ld4(tab + ((size_t)P.type2rhor[xq_type(qj) * nt1 + itype] * (P.nr + 1) + m) * 4);
          r_acc += ((cf.x * p + cf.y) * p + cf.z) * p + cf.w;
        }
      }
    }
  }
  __syncwarp();
  r_acc = group_sum<TPA>(r_acc);
  double ev[1] = {0};
  if (active && l == 0) {
    rho[i] = r_acc;
    double p = r_acc * P.rdrho + 1.0;
    int m = (int)p;
    m = max(1, min(m, P.nrho - 1));
    p -= m;
    p = fmin(p, 1.0);
    const double *cf = P.frho + ((size_t)P.type2frho[itype] * (P.nrho + 1) + m) * 8;
    const double fpi = (cf[0] * p + cf[1]) * p + cf[2];
    fp[i] = fpi;
    if (ONE) xfp[i] = make_double4(qi.x, qi.y, qi.z, fpi);
    if (EV) {
      double phi = ((cf[3] * p + cf[4]) * p + cf[5]) * p + cf[6];
      if (r_acc > P.rhomax) phi += fpi * (r_acc - P.rhomax);
      phi *= P.scale[itype * nt1 + itype];
      ev[0] = phi;
    }
  }
  if (EV) block_reduce_store<1>(ev, partials);
}

// ---- half list: pass 1 without the embedding (every listed pair adds to both atoms) ---------
// SMEM (one element, table fits): the density rows (32 bytes each) are staged in shared memory, 16-byte chunk k of row m
// at chunk m*2 + ((k + (m>>2)) & 1): the eight chunks of a 128-byte bank line then belong to (m&3, k ^ parity of m>>2), so
// a quarter warp reading chunk k of eight unrelated rows spreads over the eight 16-byte bank groups instead of piling onto
// two of them.  The block walks tiles with a grid stride so that one staging serves many tiles.
// MASK: every lane records which of its list entries were inside the force cutoff (bit `it` of a 128-bit mask per
// lane); positions do not change before the force pass of the same step, which then walks the set bits only.
template <int TPA, bool ONE, bool SMEM, bool MASK>
__global__ void __launch_bounds__(EAM_RHO_THREADS, ONE ? EAM_RHO_HALF_MINBLOCKS : (1024 / EAM_RHO_THREADS))
k_eam_rho_half(int nlocal, const double4 *__restrict__ xq, const int *__restrict__ neigh,
               const int *__restrict__ numneigh, int pitch, EamParams P, double *__restrict__ rho,
               const int *__restrict__ tile_list, int ntile_list, unsigned *__restrict__ pmask) {
  constexpr int A = 32 / TPA;
  extern __shared__ double2 s_rho2[];
  if (ONE && SMEM) {
    const double2 *g2 = reinterpret_cast<const double2 *>(P.tabC);
    const int nq = (P.nr + 1) * 2;
    for (int q = threadIdx.x; q < nq; q += blockDim.x) {
      const int m = q >> 1, k = q & 1;
      s_rho2[m * 2 + ((k + (m >> 2)) & 1)] = g2[m * 4 + k];
    }
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  const int sub = lane / TPA, l = lane % TPA;
  const int wpb = blockDim.x >> 5;
  const int nwork = tile_list ? ntile_list : (nlocal + A - 1) / A;
  const int nt1 = P.ntypes + 1;
  for (int w = blockIdx.x * wpb + (threadIdx.x >> 5); w < nwork; w += gridDim.x * wpb) {
  const int tile = tile_list ? tile_list[w] : w;
  const int i = tile * A + sub;
  const bool active = i < nlocal;
  double r_acc = 0;
  unsigned mk = 0;          // MASK: the current 32 entries of this lane; words go to pmask[tile][word][lane] as they fill
  int nit = 0;
  if (active) {
    const double4 qi = xq[i];
    const int itype = ONE ? 1 : xq_type(qi);
    const int n = numneigh[i];
    const int *nb = neigh + (size_t)tile * pitch * A + sub * TPA + l;
    const int niter = (n + TPA - 1) / TPA;
    nit = niter;
#pragma unroll kRhoUnroll
    for (int it = 0; it < niter; it++) {
      if (MASK && it && !(it & 31)) { pmask[((size_t)tile * 4 + (it >> 5) - 1) * 32 + lane] = mk; mk = 0; }
      if (it * TPA + l < n) {
        const int j = __ldg(nb + it * 32);
        const double4 qj = ld_xq(&xq[j]);
        const double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
        const double rsq = rsq_nofma(dx, dy, dz);
        if (rsq < P.cutforcesq) {
          if (MASK) mk |= 1u << (it & 31);
          double p = (rsq * rsqrt_nr(rsq)) * P.rdr + 1.0;
          int m = (int)p;
          m = min(m, P.nr - 1);
          p -= m;
          p = fmin(p, 1.0);
          if (ONE) {
            double4 cf;
            if (SMEM) {
              const int sw = (m >> 2) & 1;
              const double2 a = s_rho2[m * 2 + sw], b = s_rho2[m * 2 + (sw ^ 1)];
              cf = make_double4(a.x, a.y, b.x, b.y);
            } else cf = ld4(P.tabC + (size_t)m * 8);
            const double d = ((cf.x * p + cf.y) * p + cf.z) * p + cf.w;
            r_acc += d;
            atomicAdd(&rho[j], d);
          } else {
            const int jtype = xq_type(qj);
            const int tji = P.type2rhor[jtype * nt1 + itype], tij = P.type2rhor[itype * nt1 + jtype];
            const double4 cf = ld4(P.rhoA + ((size_t)tji * (P.nr + 1) + m) * 4);
            const double di = ((cf.x * p + cf.y) * p + cf.z) * p + cf.w;
            r_acc += di;
            double dj = di;
            if (tij != tji) {
              const double4 cg = ld4(P.rhoA + ((size_t)tij * (P.nr + 1) + m) * 4);
              dj = ((cg.x * p + cg.y) * p + cg.z) * p + cg.w;
            }
            atomicAdd(&rho[j], dj);
          }
        }
      }
    }
  }
  __syncwarp();
  r_acc = group_sum<TPA>(r_acc);
  if (active && l == 0) atomicAdd(&rho[i], r_acc);
  if (MASK) {   // the word in progress, zeros behind it (words already flushed: those below (nit-1)/32)
    const int w0 = nit ? (nit - 1) >> 5 : 0;
    for (int wd = w0; wd < 4; wd++) pmask[((size_t)tile * 4 + wd) * 32 + lane] = wd == w0 ? mk : 0u;
  }
  }
}

// embedding for the half-list path (pair_eam.cpp:231-247), one thread per owned atom
template <bool EV, bool ONE>
__global__ void __launch_bounds__(EAM_THREADS)
k_eam_embed(int nlocal, const double4 *__restrict__ xq, EamParams P, const double *__restrict__ rho,
            double *__restrict__ fp, double4 *__restrict__ xfp, double *__restrict__ partials) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int nt1 = P.ntypes + 1;
  double ev[1] = {0};
  if (i < nlocal) {
    const double4 qi = xq[i];
    const int itype = ONE ? 1 : xq_type(qi);
    const double r = rho[i];
    double p = r * P.rdrho + 1.0;
    int m = (int)p;
    m = max(1, min(m, P.nrho - 1));
    p -= m;
    p = fmin(p, 1.0);
    const double *cf = P.frho + ((size_t)P.type2frho[itype] * (P.nrho + 1) + m) * 8;
    const double fpi = (cf[0] * p + cf[1]) * p + cf[2];
    fp[i] = fpi;
    if (ONE) xfp[i] = make_double4(qi.x, qi.y, qi.z, fpi);
    if (EV) {
      double phi = ((cf[3] * p + cf[4]) * p + cf[5]) * p + cf[6];
      if (r > P.rhomax) phi += fpi * (r - P.rhomax);
      ev[0] = phi * P.scale[itype * nt1 + itype];
    }
  }
  if (EV) block_reduce_store<1>(ev, partials);
}

// ghosts of one-element runs: packed record from the ghost position and the communicated fp
__global__ void k_pack_xfp(int nghost, int nlocal, const double4 *__restrict__ xq, const double *__restrict__ fp,
                           double4 *__restrict__ xfp) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nghost) return;
  const double4 q = xq[nlocal + g];
  xfp[nlocal + g] = make_double4(q.x, q.y, q.z, fp[nlocal + g]);
}

// ---- pass 3: forces -------------------------------------------------------------------------
// MASK (half list): walk only the entries k_eam_rho_half<…,MASK> marked as inside the cutoff this step — no gather, no
// distance test and no idle lane-iterations for the ~40 % of a skin-1.0 list that lie beyond it; the next entry and its
// record are fetched while the current pair is evaluated.
template <int TPA, bool EV, bool ONE, bool SMEM, bool HALF, bool MASK>
__global__ void __launch_bounds__(EAM_THREADS, EAM_FORCE_MINBLOCKS)
k_eam_force(int nlocal, int nmax, const double4 *__restrict__ xq, const double4 *__restrict__ xfp,
            const int *__restrict__ neigh, const int *__restrict__ numneigh, int pitch, EamParams P,
            const double *__restrict__ fp, double *__restrict__ f, double *__restrict__ partials,
            const int *__restrict__ tile_list, int ntile_list, const unsigned *__restrict__ pmask) {
  constexpr int A = 32 / TPA;
  // SMEM: the 64-byte rows {density cubic, pair-term cubic} staged in shared memory, 16-byte chunk k of row m at chunk
  // m*4 + ((k + (m>>1)) & 3) (see k_eam_rho_half: a quarter warp reading chunk k of eight unrelated rows spreads over
  // the eight 16-byte groups of the banks); the block walks tiles with a grid stride, one staging serves many tiles
  extern __shared__ double2 s_tab2[];
  if (ONE && SMEM) {
    const double2 *g2 = reinterpret_cast<const double2 *>(P.tabC);
    const int nq = (P.nr + 1) * 4;
    for (int q = threadIdx.x; q < nq; q += blockDim.x) {
      const int m = q >> 2, k = q & 3;
      s_tab2[m * 4 + ((k + (m >> 1)) & 3)] = g2[q];
    }
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  const int sub = lane / TPA, l = lane % TPA;
  const int wpb = blockDim.x >> 5;
  const int nwork = tile_list ? ntile_list : (nlocal + A - 1) / A;
  const int nt1 = P.ntypes + 1;
  double ev[7] = {0, 0, 0, 0, 0, 0, 0};
  for (int w = blockIdx.x * wpb + (threadIdx.x >> 5); w < nwork; w += gridDim.x * wpb) {
  const int tile = tile_list ? tile_list[w] : w;
  const int i = tile * A + sub;
  const bool active = i < nlocal;
  double fx = 0, fy = 0, fz = 0;
  if (active) {
    const double4 qi = ONE ? xfp[i] : xq[i];
    const int itype = ONE ? 1 : xq_type(qi);
    const double fpi = ONE ? qi.w : fp[i];
    const int n = numneigh[i];
    const int *nb = neigh + (size_t)tile * pitch * A + sub * TPA + l;
    const int niter = (n + TPA - 1) / TPA;
    auto pair = [&](const int j, const double4 qj) {
        const double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
        const double rsq = rsq_nofma(dx, dy, dz);
        if (MASK || rsq < P.cutforcesq) {
          const double recip = rsqrt_nr(rsq);     // 1/r and r = rsq/r to rounding error: no IEEE sqrt / division sequences
          const double r = rsq * recip;
          double p = r * P.rdr + 1.0;
          int m = (int)p;
          m = min(m, P.nr - 1);
          p -= m;
          p = fmin(p, 1.0);
          double rhoip, rhojp, scl, fpj;
          double4 z;
          if (ONE) {
            double4 a;
            if (SMEM) {
              const int r = m >> 1;
              const double2 a0 = s_tab2[m * 4 + (r & 3)], a1 = s_tab2[m * 4 + ((r + 1) & 3)];
              const double2 z0 = s_tab2[m * 4 + ((r + 2) & 3)], z1 = s_tab2[m * 4 + ((r + 3) & 3)];
              a = make_double4(a0.x, a0.y, a1.x, a1.y);
              z = make_double4(z0.x, z0.y, z1.x, z1.y);
            } else {
              a = ld4(P.tabC + (size_t)m * 8);
              z = ld4(P.tabC + (size_t)m * 8 + 4);
            }
            rhoip = rhojp = ((3.0 * a.x * p + 2.0 * a.y) * p + a.z) * P.rdr;
            scl = P.scale11;
            fpj = qj.w;
          } else {
            const int jtype = xq_type(qj);
            const int tij = P.type2rhor[itype * nt1 + jtype], tji = P.type2rhor[jtype * nt1 + itype];
            const double4 a = ld4(P.rhoA + ((size_t)tij * (P.nr + 1) + m) * 4);
            rhoip = ((3.0 * a.x * p + 2.0 * a.y) * p + a.z) * P.rdr;
            if (tji == tij) rhojp = rhoip;
            else {
              const double4 b = ld4(P.rhoA + ((size_t)tji * (P.nr + 1) + m) * 4);
              rhojp = ((3.0 * b.x * p + 2.0 * b.y) * p + b.z) * P.rdr;
            }
            z = ld4(P.z2rV + ((size_t)P.type2z2r[itype * nt1 + jtype] * (P.nr + 1) + m) * 4);
            scl = P.scale[itype * nt1 + jtype];
            fpj = __ldg(&fp[j]);
          }
          const double z2p = ((3.0 * z.x * p + 2.0 * z.y) * p + z.z) * P.rdr;
          const double z2 = ((z.x * p + z.y) * p + z.z) * p + z.w;
          const double phi = z2 * recip;
          const double phip = z2p * recip - phi * recip;
          const double psip = fpi * rhojp + fpj * rhoip + phip;
          const double fpair = -scl * psip * recip;
          fx += dx * fpair; fy += dy * fpair; fz += dz * fpair;
          if (HALF) {
            atomicAdd(&f[j], -dx * fpair);
            atomicAdd(&f[j + nmax], -dy * fpair);
            atomicAdd(&f[j + 2 * (size_t)nmax], -dz * fpair);
          }
          if (EV) {
            ev[0] += scl * phi;
            ev[1] += dx * dx * fpair; ev[2] += dy * dy * fpair; ev[3] += dz * dz * fpair;
            ev[4] += dx * dy * fpair; ev[5] += dx * dz * fpair; ev[6] += dy * dz * fpair;
          }
        }
    };
    if (MASK) {
      const unsigned *pm = pmask + (size_t)tile * 128 + lane;
      unsigned long long m0 = pm[0] | ((unsigned long long)pm[32] << 32), m1 = pm[64] | ((unsigned long long)pm[96] << 32);
      auto next = [&]() -> int {
        if (m0) { const int b = __ffsll((long long)m0) - 1; m0 &= m0 - 1; return b; }
        if (m1) { const int b = __ffsll((long long)m1) - 1; m1 &= m1 - 1; return 64 + b; }
        return -1;
      };
      int it = next();
      int j = it >= 0 ? __ldg(nb + it * 32) : i;
      double4 qj = ld_xq(ONE ? &xfp[j] : &xq[j]);
      while (it >= 0) {
        const int itn = next();
        const int jn = itn >= 0 ? __ldg(nb + itn * 32) : i;
        const double4 qn = ld_xq(ONE ? &xfp[jn] : &xq[jn]);
        pair(j, qj);
        it = itn; j = jn; qj = qn;
      }
    } else {
#pragma unroll kForceUnroll
      for (int it = 0; it < niter; it++) {
        if (it * TPA + l < n) {
          const int j = __ldg(nb + it * 32);
          pair(j, ld_xq(ONE ? &xfp[j] : &xq[j]));
        }
      }
    }
  }
  __syncwarp();
  fx = group_sum<TPA>(fx); fy = group_sum<TPA>(fy); fz = group_sum<TPA>(fz);
  if (active && l == 0) {
    if (HALF) { atomicAdd(&f[i], fx); atomicAdd(&f[i + nmax], fy); atomicAdd(&f[i + 2 * (size_t)nmax], fz); }
    else { f[i] = fx; f[i + nmax] = fy; f[i + 2 * (size_t)nmax] = fz; }
  }
  }
  if (EV) block_reduce_store<7>(ev, partials);
}

struct EamLaunch {
  EamParams P;
  int mode;      // 0 general (multi-type), 1 one element / tables via L1, 2 one element / tables in shared memory
  size_t smem;
  bool mask;     // half list: density pass records the in-cutoff entries, force pass walks them
};

template <int TPA, bool EV, bool ONE, bool SMEM>
static int launch_pass(b200md_ctx *c, const EamLaunch &L, int pass, const int *tiles, int ntl) {
  const int A = 32 / TPA;
  const int ntiles = tiles ? ntl : (c->nlocal + A - 1) / A;
  int nblocks = (ntiles + EAM_THREADS / 32 - 1) / (EAM_THREADS / 32);
  if (nblocks == 0) return B200MD_OK;
  // shared-memory tables: resident blocks only (grid-stride over the tiles), so the table is staged once per block slot
  // (not on tally steps: the callers reduce one row of partial sums per block of the full grid)
  // Under compute/halo overlap (tile lists) the blocks must retire while the kernel runs: the pack kernels of the halo and
  // the boundary tiles sit on higher-priority streams and can only take the slots of finished blocks (a resident-blocks-only
  // grid held every SM until the interior pass ended: the fp halo arrived 280 us late) -> eight rounds of blocks there.
  if (SMEM && !EV && pass >= 1)
    nblocks = std::min(nblocks, B200_SMS * (pass == 2 ? EAM_RHO_MINBLOCKS : EAM_FORCE_MINBLOCKS) * (tiles ? 8 : 1));
  if (pass == 0) {
    auto k = k_eam_rho_fp<TPA, EV, ONE, SMEM>;
    if (SMEM) CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem));
    k<<<nblocks, EAM_THREADS, SMEM ? L.smem : 0, c->stream>>>(c->nlocal, c->xq, c->neigh, c->numneigh, c->pitch, L.P, c->rho,
                                                            c->fp, c->eam.xfp, c->partials, tiles, ntl);
  } else if (pass == 1) {
    auto k = k_eam_force<TPA, EV, ONE, SMEM, false, false>;
    if (SMEM) CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem));
    k<<<nblocks, EAM_THREADS, SMEM ? L.smem : 0, c->stream>>>(c->nlocal, c->nmax, c->xq, c->eam.xfp, c->neigh, c->numneigh,
                                                            c->pitch, L.P, c->fp, c->f, c->partials, tiles, ntl, nullptr);
  } else if (pass == 2) {   // half list, density of both atoms of a pair
    auto k = L.mask ? k_eam_rho_half<TPA, ONE, SMEM, true> : k_eam_rho_half<TPA, ONE, SMEM, false>;
    if (SMEM) CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(L.smem / 2)));
    if (!SMEM) nblocks = (ntiles + EAM_RHO_THREADS / 32 - 1) / (EAM_RHO_THREADS / 32);   // this pass writes no per-block partial sums
    k<<<nblocks, EAM_RHO_THREADS, SMEM ? L.smem / 2 : 0, c->stream>>>(c->nlocal, c->xq, c->neigh, c->numneigh, c->pitch, L.P, c->rho,
                                                                tiles, ntl, c->eam.pmask);
  } else {                  // half list, forces
    auto k = L.mask ? k_eam_force<TPA, EV, ONE, SMEM, true, true> : k_eam_force<TPA, EV, ONE, SMEM, true, false>;
    if (SMEM) CUDA_TRY(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.smem));
    k<<<nblocks, EAM_THREADS, SMEM ? L.smem : 0, c->stream>>>(
        c->nlocal, c->nmax, c->xq, c->eam.xfp, c->neigh, c->numneigh, c->pitch, L.P, c->fp, c->f, c->partials, tiles, ntl,
        c->eam.pmask);
  }
  c->launches++; c->pair_launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}
template <int TPA, bool EV>
static int launch_mode(b200md_ctx *c, const EamLaunch &L, int pass, const int *tiles, int ntl) {
  const bool force_pass = pass == 1 || pass == 3;
  if (L.mode == 2 && (force_pass || c->eam_smem > 0)) return launch_pass<TPA, EV, true, true>(c, L, pass, tiles, ntl);
  if (L.mode == 2) return launch_pass<TPA, EV, true, false>(c, L, pass, tiles, ntl);
  if (L.mode == 1) return launch_pass<TPA, EV, true, false>(c, L, pass, tiles, ntl);
  return launch_pass<TPA, EV, false, false>(c, L, pass, tiles, ntl);
}
template <int TPA>
static int launch_tpa(b200md_ctx *c, const EamLaunch &L, bool ev, int pass, const int *tiles, int ntl) {
  return ev ? launch_mode<TPA, true>(c, L, pass, tiles, ntl) : launch_mode<TPA, false>(c, L, pass, tiles, ntl);
}


// =============================================================================================
// cluster-pair list variants (atomic systems; neigh.cu k_build_cluster, pair_lj.cu k_lj_cl): L lanes share an
// i-cluster of CL = 4 atoms kept in registers, each j record is gathered once per cluster, and rho_j / f_j leave
// with one RED (triple) per (cluster, j).  Table rows are still fetched per evaluated pair.
// =============================================================================================
#ifndef EAM_CL_MINBLOCKS
#define EAM_CL_MINBLOCKS 2
#endif

// pass 1.  HALF: rho_i and rho_j of every listed pair (RED), embedding done by k_eam_embed after the reverse_comm.
// Full list: rho_i only, then the embedding of the cluster's atoms (the reference Kokkos port's KernelAB).
template <int L, bool EV, bool ONE, bool HALF>
__global__ void __launch_bounds__(EAM_THREADS, EAM_CL_MINBLOCKS)
k_eam_rho_cl(int nlocal, int ncl, const double4 *__restrict__ xq, const int *__restrict__ neigh, const int *__restrict__ numneigh,
             int pitch, EamParams P, double *__restrict__ rho, double *__restrict__ fp, double4 *__restrict__ xfp,
             double *__restrict__ partials, const int *__restrict__ tile_list, int ntile_list) {
  constexpr int C = 32 / L;
  const int lane = threadIdx.x & 31;
  const int sub = lane / L, l = lane % L;
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const bool in_list = !tile_list || w < ntile_list;
  const int tile = tile_list ? (in_list ? tile_list[w] : 0) : w;
  const int c = tile * C + sub;
  const bool active = in_list && c < ncl;
  const int i0 = c * CL;
  const int nt1 = P.ntypes + 1;
  double xi[CL], yi[CL], zi[CL], ri[CL];
  int ti[CL];
#pragma unroll
  for (int k = 0; k < CL; k++) { xi[k] = yi[k] = zi[k] = ri[k] = 0; ti[k] = 1; }
  if (active) {
#pragma unroll
    for (int k = 0; k < CL; k++) {
      const double4 q = xq[min(i0 + k, nlocal - 1)];
      xi[k] = q.x; yi[k] = q.y; zi[k] = q.z;
      if (!ONE) ti[k] = xq_type(q);
    }
    const int n = numneigh[c];
    const int *nb = neigh + (size_t)tile * pitch * C + sub * L + l;
    const int niter = (n + L - 1) / L;
    // software pipeline as in k_lj_cl: entry it+2 and record it+1 in flight while slot it is evaluated; the four pairs
    // of a slot are evaluated branch-free (a miss runs at r = 1 with weight 0) so their table loads and FP64 chains overlap
    unsigned e0 = (0 * L + l < n) ? (unsigned)__ldg(nb) : 0u;
    unsigned e1 = (1 * L + l < n) ? (unsigned)__ldg(nb + 32) : 0u;
    double4 q0 = ld_xq(&xq[e0 & CL_JMASK]);
    for (int it = 0; it < niter; it++) {
      const unsigned e2 = ((it + 2) * L + l < n) ? (unsigned)__ldg(nb + (it + 2) * 32) : 0u;
      const double4 q1 = ld_xq(&xq[e1 & CL_JMASK]);
      {
        const int j = (int)(e0 & CL_JMASK);
        const unsigned m4 = e0 >> CL_SHIFT;
        const double4 qj = q0;
        const int tj = ONE ? 1 : xq_type(qj);
        double g = 0;
#pragma unroll
        for (int k = 0; k < CL; k++) {
          const double dx = xi[k] - qj.x, dy = yi[k] - qj.y, dz = zi[k] - qj.z;
          const double rsq = fma(dz, dz, fma(dy, dy, dx * dx));
          const bool hit = ((m4 >> k) & 1u) && rsq < P.cutforcesq;
          const double rs = hit ? rsq : 1.0;
          double p = rs * rsqrt_nr(rs) * P.rdr + 1.0;
          int m = (int)p;
          m = min(m, P.nr - 1);
          p -= m;
          p = fmin(p, 1.0);
          if (ONE) {
            const double4 cf = ld4_if(P.tabC + (size_t)m * 8, hit);
            const double d = hit ? ((cf.x * p + cf.y) * p + cf.z) * p + cf.w : 0.0;
            ri[k] += d;
            if (HALF) g += d;
          } else {
            const int tji = P.type2rhor[tj * nt1 + ti[k]], tij = P.type2rhor[ti[k] * nt1 + tj];
            const double4 cf = ld4(P.rhoA + ((size_t)tji * (P.nr + 1) + m) * 4);
            const double di = hit ? ((cf.x * p + cf.y) * p + cf.z) * p + cf.w : 0.0;
            ri[k] += di;
            if (HALF) {
              double dj = di;
              if (tij != tji) {
                const double4 cg = ld4(P.rhoA + ((size_t)tij * (P.nr + 1) + m) * 4);
                dj = hit ? ((cg.x * p + cg.y) * p + cg.z) * p + cg.w : 0.0;
              }
              g += dj;
            }
          }
        }
        if (HALF && g != 0.0) atomicAdd(&rho[j], g);
      }
      e0 = e1; e1 = e2; q0 = q1;
    }
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < CL; k++) ri[k] = group_sum<L>(ri[k]);
  double ev[1] = {0};
  if (active) {
#pragma unroll
    for (int k = 0; k < CL; k++) {
      if (l == (L >= CL ? k : 0) && i0 + k < nlocal) {
        const int i = i0 + k;
        if (HALF) atomicAdd(&rho[i], ri[k]);
        else {
          const double r = ri[k];
          rho[i] = r;
          double p = r * P.rdrho + 1.0;
          int m = (int)p;
          m = max(1, min(m, P.nrho - 1));
          p -= m;
          p = fmin(p, 1.0);
          const double *cf = P.frho + ((size_t)P.type2frho[ti[k]] * (P.nrho + 1) + m) * 8;
          const double fpi = (cf[0] * p + cf[1]) * p + cf[2];
          fp[i] = fpi;
          if (ONE) xfp[i] = make_double4(xi[k], yi[k], zi[k], fpi);
          if (EV) {
            double phi = ((cf[3] * p + cf[4]) * p + cf[5]) * p + cf[6];
            if (r > P.rhomax) phi += fpi * (r - P.rhomax);
            ev[0] += phi * P.scale[ti[k] * nt1 + ti[k]];
          }
        }
      }
    }
  }
  if (EV && !HALF) block_reduce_store<1>(ev, partials);
}

// pass 3: forces
template <int L, bool EV, bool ONE, bool HALF>
__global__ void __launch_bounds__(EAM_THREADS, EAM_CL_MINBLOCKS)
k_eam_force_cl(int nlocal, int ncl, int nmax, const double4 *__restrict__ xq, const double4 *__restrict__ xfp,
               const int *__restrict__ neigh, const int *__restrict__ numneigh, int pitch, EamParams P,
               const double *__restrict__ fp, double *__restrict__ f, double *__restrict__ partials,
               const int *__restrict__ tile_list, int ntile_list) {
  constexpr int C = 32 / L;
  const int lane = threadIdx.x & 31;
  const int sub = lane / L, l = lane % L;
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const bool in_list = !tile_list || w < ntile_list;
  const int tile = tile_list ? (in_list ? tile_list[w] : 0) : w;
  const int c = tile * C + sub;
  const bool active = in_list && c < ncl;
  const int i0 = c * CL;
  const int nt1 = P.ntypes + 1;
  double xi[CL], yi[CL], zi[CL], fpi[CL], fx[CL], fy[CL], fz[CL];
  int ti[CL];
  double ev[7] = {0, 0, 0, 0, 0, 0, 0};
#pragma unroll
  for (int k = 0; k < CL; k++) { xi[k] = yi[k] = zi[k] = fpi[k] = fx[k] = fy[k] = fz[k] = 0; ti[k] = 1; }
  if (active) {
#pragma unroll
    for (int k = 0; k < CL; k++) {
      const int i = min(i0 + k, nlocal - 1);
      const double4 q = ONE ? xfp[i] : xq[i];
      xi[k] = q.x; yi[k] = q.y; zi[k] = q.z;
      if (ONE) fpi[k] = q.w; else { ti[k] = xq_type(q); fpi[k] = fp[i]; }
    }
    const int n = numneigh[c];
    const int *nb = neigh + (size_t)tile * pitch * C + sub * L + l;
    const int niter = (n + L - 1) / L;
    const double4 *rec = ONE ? xfp : xq;
    unsigned e0 = (0 * L + l < n) ? (unsigned)__ldg(nb) : 0u;
    unsigned e1 = (1 * L + l < n) ? (unsigned)__ldg(nb + 32) : 0u;
    double4 q0 = ld_xq(&rec[e0 & CL_JMASK]);
    for (int it = 0; it < niter; it++) {
      const unsigned e2 = ((it + 2) * L + l < n) ? (unsigned)__ldg(nb + (it + 2) * 32) : 0u;
      const double4 q1 = ld_xq(&rec[e1 & CL_JMASK]);
      {
        const int j = (int)(e0 & CL_JMASK);
        const unsigned m4 = e0 >> CL_SHIFT;
        const double4 qj = q0;
        const int tj = ONE ? 1 : xq_type(qj);
        const double fpj = ONE ? qj.w : __ldg(&fp[j]);
        double gx = 0, gy = 0, gz = 0;
#pragma unroll
        for (int k = 0; k < CL; k++) {
          const double dx = xi[k] - qj.x, dy = yi[k] - qj.y, dz = zi[k] - qj.z;
          const double rsq = fma(dz, dz, fma(dy, dy, dx * dx));
          const bool hit = ((m4 >> k) & 1u) && rsq < P.cutforcesq;
          const double rs = hit ? rsq : 1.0;
          const double recip = rsqrt_nr(rs);
          const double r = rs * recip;
          double p = r * P.rdr + 1.0;
          int m = (int)p;
          m = min(m, P.nr - 1);
          p -= m;
          p = fmin(p, 1.0);
          double rhoip, rhojp, scl;
          double4 z;
          if (ONE) {
            const double4 a = ld4_if(P.tabC + (size_t)m * 8, hit);
            z = ld4_if(P.tabC + (size_t)m * 8 + 4, hit);
            rhoip = rhojp = ((3.0 * a.x * p + 2.0 * a.y) * p + a.z) * P.rdr;
            scl = P.scale11;
          } else {
            const int tij = P.type2rhor[ti[k] * nt1 + tj], tji = P.type2rhor[tj * nt1 + ti[k]];
            const double4 a = ld4(P.rhoA + ((size_t)tij * (P.nr + 1) + m) * 4);
            rhoip = ((3.0 * a.x * p + 2.0 * a.y) * p + a.z) * P.rdr;
            if (tji == tij) rhojp = rhoip;
            else {
              const double4 b = ld4(P.rhoA + ((size_t)tji * (P.nr + 1) + m) * 4);
              rhojp = ((3.0 * b.x * p + 2.0 * b.y) * p + b.z) * P.rdr;
            }
            z = ld4(P.z2rV + ((size_t)P.type2z2r[ti[k] * nt1 + tj] * (P.nr + 1) + m) * 4);
            scl = P.scale[ti[k] * nt1 + tj];
          }
          const double z2p = ((3.0 * z.x * p + 2.0 * z.y) * p + z.z) * P.rdr;
          const double z2 = ((z.x * p + z.y) * p + z.z) * p + z.w;
          const double phi = z2 * recip;
          const double phip = z2p * recip - phi * recip;
          const double psip = fpi[k] * rhojp + fpj * rhoip + phip;
          const double fpair = hit ? -scl * psip * recip : 0.0;
          fx[k] = fma(dx, fpair, fx[k]); fy[k] = fma(dy, fpair, fy[k]); fz[k] = fma(dz, fpair, fz[k]);
          if (HALF) { gx = fma(dx, fpair, gx); gy = fma(dy, fpair, gy); gz = fma(dz, fpair, gz); }
          if (EV) {
            const double px = dx * fpair, py = dy * fpair, pz = dz * fpair;
            ev[0] += hit ? scl * phi : 0.0;
            ev[1] += dx * px; ev[2] += dy * py; ev[3] += dz * pz;
            ev[4] += dx * py; ev[5] += dx * pz; ev[6] += dy * pz;
          }
        }
        if (HALF && m4 && (gx != 0.0 || gy != 0.0 || gz != 0.0)) {
          atomicAdd(&f[j], -gx);
          atomicAdd(&f[j + nmax], -gy);
          atomicAdd(&f[j + 2 * (size_t)nmax], -gz);
        }
      }
      e0 = e1; e1 = e2; q0 = q1;
    }
  }
  __syncwarp();
#pragma unroll
  for (int k = 0; k < CL; k++) { fx[k] = group_sum<L>(fx[k]); fy[k] = group_sum<L>(fy[k]); fz[k] = group_sum<L>(fz[k]); }
  if (active) {
#pragma unroll
    for (int k = 0; k < CL; k++) {
      if (l == (L >= CL ? k : 0) && i0 + k < nlocal) {
        const int i = i0 + k;
        if (HALF) { atomicAdd(&f[i], fx[k]); atomicAdd(&f[i + nmax], fy[k]); atomicAdd(&f[i + 2 * (size_t)nmax], fz[k]); }
        else { f[i] = fx[k]; f[i + nmax] = fy[k]; f[i + 2 * (size_t)nmax] = fz[k]; }
      }
    }
  }
  if (EV) block_reduce_store<7>(ev, partials);
}

template <int L, bool EV, bool ONE>
static int launch_pass_cl(b200md_ctx *c, const EamParams &P, int pass, const int *tiles, int ntl) {
  const int ncl = list_units(c);
  const int ntiles = tiles ? ntl : list_tiles(c);
  const int nblocks = (ntiles + EAM_THREADS / 32 - 1) / (EAM_THREADS / 32);
  if (nblocks == 0) return B200MD_OK;
  if (pass == 0)
    k_eam_rho_cl<L, EV, ONE, false><<<nblocks, EAM_THREADS, 0, c->stream>>>(c->nlocal, ncl, c->xq, c->neigh, c->numneigh, c->pitch, P,
                                                                          c->rho, c->fp, c->eam.xfp, c->partials, tiles, ntl);
  else if (pass == 1)
    k_eam_force_cl<L, EV, ONE, false><<<nblocks, EAM_THREADS, 0, c->stream>>>(c->nlocal, ncl, c->nmax, c->xq, c->eam.xfp, c->neigh,
                                                                            c->numneigh, c->pitch, P, c->fp, c->f, c->partials, tiles, ntl);
  else if (pass == 2)
    k_eam_rho_cl<L, false, ONE, true><<<nblocks, EAM_THREADS, 0, c->stream>>>(c->nlocal, ncl, c->xq, c->neigh, c->numneigh, c->pitch, P,
                                                                            c->rho, c->fp, c->eam.xfp, c->partials, tiles, ntl);
  else
    k_eam_force_cl<L, EV, ONE, true><<<nblocks, EAM_THREADS, 0, c->stream>>>(c->nlocal, ncl, c->nmax, c->xq, c->eam.xfp, c->neigh,
                                                                           c->numneigh, c->pitch, P, c->fp, c->f, c->partials, tiles, ntl);
  c->launches++; c->pair_launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}
template <int L>
static int launch_cl(b200md_ctx *c, const EamLaunch &Ln, bool ev, int pass, const int *tiles, int ntl) {
  if (Ln.mode) return ev ? launch_pass_cl<L, true, true>(c, Ln.P, pass, tiles, ntl) : launch_pass_cl<L, false, true>(c, Ln.P, pass, tiles, ntl);
  return ev ? launch_pass_cl<L, true, false>(c, Ln.P, pass, tiles, ntl) : launch_pass_cl<L, false, false>(c, Ln.P, pass, tiles, ntl);
}

EamParams pair_eam_params(const b200md_ctx *c) {
  const EamTables &e = c->eam;
  return {e.ntypes, e.nrho, e.nr, e.rdr, e.rdrho, e.rhomax, e.cutforcesq, e.scale11, e.frho, e.rhoA, e.z2rV,
          e.tabC, e.scale, e.type2frho, e.type2rhor, e.type2z2r};
}

static int eam_prepare(b200md_ctx *c, EamLaunch &L) {
  EamTables &e = c->eam;
  if (e.one && e.xfp_cap < c->nmax) {
    if (e.xfp) CUDA_TRY(cudaFree(e.xfp));
    CUDA_TRY(cudaMalloc(&e.xfp, sizeof(double4) * (size_t)c->nmax));
    e.xfp_cap = c->nmax;
  }
  L.P = pair_eam_params(c);
  // in-cutoff masks (half list, per-atom rows): 128 bits per lane, one lane's share of a row must fit
  L.mask = c->eam_mask && c->list_kind == B200MD_LIST_HALF_NEWTON && !c->cluster && c->pitch / c->tpa <= 128;
  if (L.mask) {
    const size_t need = ((size_t)c->nmax / (32 / c->tpa) + 2) * 32;
    if (need > e.pmask_cap) {
      if (e.pmask) CUDA_TRY(cudaFree(e.pmask));
      CUDA_TRY(cudaMalloc(&e.pmask, sizeof(unsigned) * 4 * need));
      e.pmask_cap = need;
    }
  }
  L.smem = (size_t)(e.nr + 1) * 8 * sizeof(double);
  // measured on B200 (2.05 M atoms, Cu_u3, half list): rows copied as they are into shared memory pile the reads of a
  // quarter warp onto two 16-byte bank groups (round 1: 681 against 752 Matom-steps/s through L1); with the rows
  // swizzled the force pass gains 7 % (815 -> 756 us), the density pass nothing (460 -> 462 us).  auto: force pass only.
  L.mode = 0;
  if (e.one) L.mode = (L.smem <= 64 * 1024 && c->eam_smem != 0) ? 2 : 1;
  return B200MD_OK;
}

// pass: 0 = density + embedding (k_eam_rho_fp), 1 = forces (k_eam_force); half list: 2 = density, 3 = forces
int pair_eam_launch(b200md_ctx *c, bool ev, int pass, const int *tiles, int ntl) {
  EamLaunch L;
  TRY(eam_prepare(c, L));
  if (c->cluster) switch (c->tpa) {
    case 1: return launch_cl<1>(c, L, ev, pass, tiles, ntl);
    case 2: return launch_cl<2>(c, L, ev, pass, tiles, ntl);
    case 4: return launch_cl<4>(c, L, ev, pass, tiles, ntl);
    case 8: return launch_cl<8>(c, L, ev, pass, tiles, ntl);
    case 16: return launch_cl<16>(c, L, ev, pass, tiles, ntl);
    case 32: return launch_cl<32>(c, L, ev, pass, tiles, ntl);
  }
  else switch (c->tpa) {
    case 1: return launch_tpa<1>(c, L, ev, pass, tiles, ntl);
    case 2: return launch_tpa<2>(c, L, ev, pass, tiles, ntl);
    case 4: return launch_tpa<4>(c, L, ev, pass, tiles, ntl);
    case 8: return launch_tpa<8>(c, L, ev, pass, tiles, ntl);
    case 16: return launch_tpa<16>(c, L, ev, pass, tiles, ntl);
    case 32: return launch_tpa<32>(c, L, ev, pass, tiles, ntl);
  }
  b200md_set_error("b200md: lanes per atom must be 1, 2, 4, 8, 16 or 32");
  return B200MD_ERR;
}

// PairEAM::compute's forward_comm(this): fp of owners -> ghosts (pair_eam.cpp:251, pack :855-876)
// one rank, one element: fp of the owner and the packed ghost record in one pass
__global__ void k_fp_xfp_flat(int nghost, int nlocal, const double4 *__restrict__ xq, double *__restrict__ fp,
                              const int *__restrict__ ghost_src, double4 *__restrict__ xfp) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nghost) return;
  const double v = fp[ghost_src[nlocal + g]];
  const double4 q = xq[nlocal + g];
  fp[nlocal + g] = v;
  xfp[nlocal + g] = make_double4(q.x, q.y, q.z, v);
}

int pair_eam_fp_comm(b200md_ctx *c) {
  if (c->all_self && c->eam.one) {
    if (c->nghost > 0) {
      k_fp_xfp_flat<<<(c->nghost + 255) / 256, 256, 0, c->stream>>>(c->nghost, c->nlocal, c->xq, c->fp, c->ghost_src, c->eam.xfp);
      c->launches++;
      CUDA_TRY(cudaGetLastError());
    }
    return B200MD_OK;
  }
  TRY(comm_forward_scalar(c, c->fp));
  if (c->eam.one && c->nghost > 0) {
    k_pack_xfp<<<(c->nghost + 255) / 256, 256, 0, c->stream>>>(c->nghost, c->nlocal, c->xq, c->fp, c->eam.xfp);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
  }
  return B200MD_OK;
}

int pair_eam_embed(b200md_ctx *c, bool ev, double *e_embed) {
  const int nb_atoms = (c->nlocal + EAM_THREADS - 1) / EAM_THREADS;
  if (!nb_atoms) return B200MD_OK;
  EamLaunch L;
  TRY(eam_prepare(c, L));
  if (ev) TRY(ensure_partials(c, nb_atoms));
  if (ev) { if (L.mode) k_eam_embed<true, true><<<nb_atoms, EAM_THREADS, 0, c->stream>>>(c->nlocal, c->xq, L.P, c->rho, c->fp, c->eam.xfp, c->partials);
            else k_eam_embed<true, false><<<nb_atoms, EAM_THREADS, 0, c->stream>>>(c->nlocal, c->xq, L.P, c->rho, c->fp, c->eam.xfp, c->partials); }
  else { if (L.mode) k_eam_embed<false, true><<<nb_atoms, EAM_THREADS, 0, c->stream>>>(c->nlocal, c->xq, L.P, c->rho, c->fp, c->eam.xfp, c->partials);
         else k_eam_embed<false, false><<<nb_atoms, EAM_THREADS, 0, c->stream>>>(c->nlocal, c->xq, L.P, c->rho, c->fp, c->eam.xfp, c->partials); }
  c->launches++; c->pair_launches++;
  CUDA_TRY(cudaGetLastError());
  if (ev && e_embed) { double e1[1]; TRY(reduce_partials(c, nb_atoms, 1, e1)); *e_embed = e1[0]; }
  return B200MD_OK;
}

// half list / newton on: PairEAM::compute with its reverse_comm(rho) and, via the caller,
// Comm::reverse_comm of the forces
static int pair_eam_compute_half(b200md_ctx *c, bool ev) {
  const int ntiles = list_tiles(c);
  const int nblocks = (ntiles + EAM_THREADS / 32 - 1) / (EAM_THREADS / 32);
  const int nall = c->nlocal + c->nghost;
  const int nb_atoms = (c->nlocal + EAM_THREADS - 1) / EAM_THREADS;
  c->ev_valid = false;
  CUDA_TRY(cudaMemsetAsync(c->rho, 0, sizeof(double) * nall, c->stream));
  CUDA_TRY(cudaMemset2DAsync(c->f, sizeof(double) * (size_t)c->nmax, 0, sizeof(double) * nall, 3, c->stream));
  double e_embed = 0, out7[7] = {0, 0, 0, 0, 0, 0, 0};
  if (nblocks) {
    if (ev) TRY(ensure_partials(c, nblocks > nb_atoms ? nblocks : nb_atoms));
    TRY(pair_eam_launch(c, false, 2, nullptr, 0));
  }
  TRY(comm_reverse_scalar(c, c->rho));
  TRY(pair_eam_embed(c, ev, &e_embed));
  TRY(pair_eam_fp_comm(c));
  if (nblocks) {
    TRY(pair_eam_launch(c, ev, 3, nullptr, 0));
    if (ev) TRY(reduce_partials(c, nblocks, 7, out7));
  }
  if (ev) {
    c->eng_vdwl = e_embed + out7[0];
    for (int k = 0; k < 6; k++) c->virial[k] = out7[1 + k];
    c->ev_valid = true;
  }
  return B200MD_OK;
}

int pair_eam_compute(b200md_ctx *c, int eflag, int vflag) {
  if (c->list_kind == B200MD_LIST_HALF_NEWTON) return pair_eam_compute_half(c, eflag || vflag);
  const bool ev = eflag || vflag;
  const int ntiles = list_tiles(c);
  const int nblocks = (ntiles + EAM_THREADS / 32 - 1) / (EAM_THREADS / 32);
  c->ev_valid = false;
  if (ev && nblocks) TRY(ensure_partials(c, nblocks));
  double e_embed = 0, out7[7] = {0, 0, 0, 0, 0, 0, 0};
  TRY(pair_eam_launch(c, ev, 0, nullptr, 0));
  if (ev && nblocks) { double e1[1]; TRY(reduce_partials(c, nblocks, 1, e1)); e_embed = e1[0]; }
  TRY(pair_eam_fp_comm(c));
  TRY(pair_eam_launch(c, ev, 1, nullptr, 0));
  if (ev && nblocks) TRY(reduce_partials(c, nblocks, 7, out7));
  if (ev) {
    c->eng_vdwl = e_embed + 0.5 * out7[0];
    for (int k = 0; k < 6; k++) c->virial[k] = 0.5 * out7[1 + k];
    c->ev_valid = true;
  }
  return B200MD_OK;
}
