// pair_lj.cu — pair_style lj/cut force kernel.
//
// Reference: PairLJCut::compute (/root/reference/src/pair_lj_cut.cpp:69-137):
//   r2inv = 1/rsq; r6inv = r2inv^3; forcelj = r6inv*(lj1*r6inv - lj2); fpair = forcelj*r2inv
//   evdwl = r6inv*(lj3*r6inv - lj4) - offset          (rsq < cutsq, strict)
// with Verlet::force_clear (src/verlet.cpp:376-421) fused in (the full-list kernel writes
// f[i] instead of accumulating; reference Kokkos does the same with ZEROFLAG,
// src/KOKKOS/pair_kokkos.h:141-145), and the energy/virial tallies of Pair::ev_tally
// (src/pair.cpp:1013-1108) done as warp- then block-level reductions, no atomics.
//
// Full list (newton off): TPA lanes cooperate on atom i, each lane takes every TPA-th
// neighbor from the warp-tile-transposed list (a warp reads 128 contiguous bytes of indices
// per iteration), gathers the 32-byte {x,y,z,type} record of j, and the partial forces are
// shuffle-reduced; lane 0 writes f[i].  Each pair is visited from both sides, so energy and
// virial take weight 1/2.  Half list (newton on): same traversal, f[j] -= via FP64 RED.
#include "ctx.h"
#include "util.cuh"

#define PAIR_THREADS 256
#define LJ_MAX_TYPES_SMEM 8
// tuning macros (profiles/r01_summary.md has the sweep on B200): neighbors in flight per lane and
// minimum resident blocks per SM (a software prefetch of the next index batch was tried and lost
// to plain occupancy).  The kernel
// is latency-bound (ncu: long-scoreboard stalls, no unit above 70 %), and thread-level
// parallelism beats unrolling: 32 registers / 8 blocks per SM / no unroll is +23 % over
// 48 registers / unroll 4.  The tally variants (thermo steps only) keep 4 blocks per SM.
#ifndef LJ_UNROLL
#define LJ_UNROLL 1
#endif
#ifndef LJ_MINBLOCKS
#define LJ_MINBLOCKS 8
#endif

struct LJOne { double cutsq, lj1, lj2, lj3, lj4, offset; };

template <int TPA, bool EV, bool HALF, bool ONETYPE>
__global__ void __launch_bounds__(PAIR_THREADS, (EV || !ONETYPE) ? 4 : LJ_MINBLOCKS)
k_lj(int nlocal, int nmax, const double4 *__restrict__ xq, const int *__restrict__ neigh,
     const int *__restrict__ numneigh, int pitch, LJOne one, const double *__restrict__ ljtab, int ntypes,
     double *__restrict__ f, double *__restrict__ partials, const int *__restrict__ tile_list, int ntile_list) {
  constexpr int A = 32 / TPA;
  constexpr int kUnroll = LJ_UNROLL;
  __shared__ double s_tab[ONETYPE ? 1 : (LJ_MAX_TYPES_SMEM + 1) * (LJ_MAX_TYPES_SMEM + 1) * 6];
  const int nt1 = ntypes + 1;
  if (!ONETYPE) {
    for (int r = threadIdx.x; r < nt1 * nt1; r += blockDim.x)
#pragma unroll
      for (int k = 0; k < 6; k++) s_tab[r * 6 + k] = ljtab[r * 8 + k];
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  const int sub = lane / TPA, l = lane % TPA;
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const bool in_list = !tile_list || w < ntile_list;
  const int tile = tile_list ? (in_list ? tile_list[w] : 0) : w;
  const int i = tile * A + sub;
  const bool active = in_list && i < nlocal;

  double fx = 0, fy = 0, fz = 0;
  double ev[7] = {0, 0, 0, 0, 0, 0, 0};  // evdwl, vxx, vyy, vzz, vxy, vxz, vyz
  if (active) {
    const double4 qi = xq[i];
    const int n = numneigh[i];
    const int itype = ONETYPE ? 1 : xq_type(qi);
    const int *nb = neigh + (size_t)tile * pitch * A + sub * TPA + l;
    const int niter = (n + TPA - 1) / TPA;
#pragma unroll kUnroll
    for (int it = 0; it < niter; it++) {
      if (it * TPA + l < n) {
        const int j = __ldg(nb + it * 32);
        const double4 qj = ld_xq(&xq[j]);
        const double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
        const double rsq = rsq_nofma(dx, dy, dz);
        double cutsq, lj1, lj2, lj3, lj4, offset;
        if (ONETYPE) { cutsq = one.cutsq; lj1 = one.lj1; lj2 = one.lj2; lj3 = one.lj3; lj4 = one.lj4; offset = one.offset; }
        else {
          const double *t = &s_tab[(itype * nt1 + xq_type(qj)) * 6];
          cutsq = t[0]; lj1 = t[1]; lj2 = t[2]; lj3 = t[3]; lj4 = t[4]; offset = t[5];
        }
        if (rsq < cutsq) {
          const double r2inv = 1.0 / rsq;
          const double r6inv = r2inv * r2inv * r2inv;
          const double forcelj = r6inv * (lj1 * r6inv - lj2);
          const double fpair = forcelj * r2inv;
          fx += dx * fpair; fy += dy * fpair; fz += dz * fpair;
          if (HALF) {
            atomicAdd(&f[j], -dx * fpair);
            atomicAdd(&f[j + nmax], -dy * fpair);
            atomicAdd(&f[j + 2 * (size_t)nmax], -dz * fpair);
          }
          if (EV) {
            ev[0] += r6inv * (lj3 * r6inv - lj4) - offset;
            ev[1] += dx * dx * fpair; ev[2] += dy * dy * fpair; ev[3] += dz * dz * fpair;
            ev[4] += dx * dy * fpair; ev[5] += dx * dz * fpair; ev[6] += dy * dz * fpair;
          }
        }
      }
    }
  }
  __syncwarp();
  fx = group_sum<TPA>(fx); fy = group_sum<TPA>(fy); fz = group_sum<TPA>(fz);
  if (active && l == 0) {
    if (HALF) {
      atomicAdd(&f[i], fx); atomicAdd(&f[i + nmax], fy); atomicAdd(&f[i + 2 * (size_t)nmax], fz);
    } else {
      f[i] = fx; f[i + nmax] = fy; f[i + 2 * (size_t)nmax] = fz;
    }
  }
  if (EV) block_reduce_store<7>(ev, partials);
}

template <int TPA, bool EV, bool HALF>
static void launch_lj2(b200md_ctx *c, int nblocks, const LJOne &one, const int *tiles, int ntl) {
  if (c->ntypes == 1)
    k_lj<TPA, EV, HALF, true><<<nblocks, PAIR_THREADS, 0, c->stream>>>(c->nlocal, c->nmax, c->xq, c->neigh, c->numneigh,
                                                                      c->pitch, one, c->ljtab, c->ntypes, c->f, c->partials, tiles, ntl);
  else
    k_lj<TPA, EV, HALF, false><<<nblocks, PAIR_THREADS, 0, c->stream>>>(c->nlocal, c->nmax, c->xq, c->neigh, c->numneigh,
                                                                       c->pitch, one, c->ljtab, c->ntypes, c->f, c->partials, tiles, ntl);
}
template <int TPA>
static void launch_lj(b200md_ctx *c, int nblocks, bool ev, const LJOne &one, const int *tiles, int ntl) {
  const bool half = c->list_kind == B200MD_LIST_HALF_NEWTON;
  if (ev) { if (half) launch_lj2<TPA, true, true>(c, nblocks, one, tiles, ntl); else launch_lj2<TPA, true, false>(c, nblocks, one, tiles, ntl); }
  else { if (half) launch_lj2<TPA, false, true>(c, nblocks, one, tiles, ntl); else launch_lj2<TPA, false, false>(c, nblocks, one, tiles, ntl); }
}


// ---- cluster-pair list (atomic systems; neigh.cu k_build_cluster) ---------------------------------------
// L lanes share one i-cluster of CL = 4 owned atoms whose positions and force accumulators live in
// registers; each lane takes every L-th entry of the cluster's j-list, gathers x_j ONCE, evaluates the pairs
// its 4-bit mask names, and (half list) returns the summed f_j with ONE RED triple.  Against the per-atom
// list this cuts the gathers and REDs per atom ~2.2x (the L1 data pipe is what binds these kernels) at
// the price of distance tests on masked-out slots (FP64 pipe was at 26 %).
// 512 threads per SM (117 registers) as blocks of LJ_CL_THREADS: measured at 2.05 M atoms, pair time 0.406 ms with 2 blocks
// of 256 threads, 0.390 with 4 x 128, 0.372 with 8 x 64, 0.364 with 16 x 32.  The warps of a block share nothing (the
// one-type constants are kernel arguments), but a block's registers stay allocated until its slowest warp has walked its
// clusters' lists: one warp per block returns every slot as soon as its own work ends.
#ifndef LJ_CL_THREADS
#define LJ_CL_THREADS 32
#endif
#ifndef LJ_CL_MINBLOCKS
#define LJ_CL_MINBLOCKS (512 / LJ_CL_THREADS)
#endif
template <int L, bool EV, bool HALF, bool ONETYPE>
__global__ void __launch_bounds__(LJ_CL_THREADS, LJ_CL_MINBLOCKS)
k_lj_cl(int nlocal, int ncl, int nmax, const double4 *__restrict__ xq, const int *__restrict__ neigh,
        const int *__restrict__ numneigh, int pitch, LJOne one, const double *__restrict__ ljtab, int ntypes,
        double *__restrict__ f, double *__restrict__ partials, const int *__restrict__ tile_list, int ntile_list) {
  constexpr int C = 32 / L;
  __shared__ double s_tab[ONETYPE ? 1 : (LJ_MAX_TYPES_SMEM + 1) * (LJ_MAX_TYPES_SMEM + 1) * 6];
  const int nt1 = ntypes + 1;
  if (!ONETYPE) {
    for (int r = threadIdx.x; r < nt1 * nt1; r += blockDim.x)
#pragma unroll
      for (int k = 0; k < 6; k++) s_tab[r * 6 + k] = ljtab[r * 8 + k];
    __syncthreads();
  }
  const int lane = threadIdx.x & 31;
  const int sub = lane / L, l = lane % L;
  const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const bool in_list = !tile_list || w < ntile_list;
  const int tile = tile_list ? (in_list ? tile_list[w] : 0) : w;
  const int c = tile * C + sub;
  const bool active = in_list && c < ncl;
  const int i0 = c * CL;

  double xi[CL], yi[CL], zi[CL], fx[CL], fy[CL], fz[CL];
  int ti[CL];
  double ev[7] = {0, 0, 0, 0, 0, 0, 0};  // evdwl, vxx, vyy, vzz, vxy, vxz, vyz
#pragma unroll
  for (int k = 0; k < CL; k++) { fx[k] = fy[k] = fz[k] = 0; xi[k] = yi[k] = zi[k] = 0; ti[k] = 1; }
  if (active) {
#pragma unroll
    for (int k = 0; k < CL; k++) {
      const double4 q = xq[min(i0 + k, nlocal - 1)];   // absent atoms of the last cluster never appear in a mask
      xi[k] = q.x; yi[k] = q.y; zi[k] = q.z;
      if (!ONETYPE) ti[k] = xq_type(q);
    }
    const int n = numneigh[c];
    const int *nb = neigh + (size_t)tile * pitch * C + sub * L + l;
    const int niter = (n + L - 1) / L;
    // software pipeline (the kernel is latency-bound: ncu long-scoreboard stalls): the entry of iteration it+2 and
    // the x_j record of iteration it+1 are in flight while iteration it is evaluated.  Slots past the end of the
    // list read entry 0 of the padded row with an empty mask.
    unsigned e0 = (0 * L + l < n) ? (unsigned)__ldg(nb) : 0u;
    unsigned e1 = (1 * L + l < n) ? (unsigned)__ldg(nb + 32) : 0u;
    double4 q0 = ld_xq(&xq[e0 & CL_JMASK]);
    for (int it = 0; it < niter; it++) {
      const unsigned e2 = ((it + 2) * L + l < n) ? (unsigned)__ldg(nb + (it + 2) * 32) : 0u;
      const double4 q1 = ld_xq(&xq[e1 & CL_JMASK]);
      {
        const int j = (int)(e0 & CL_JMASK);
        const unsigned m = e0 >> CL_SHIFT;
        const double4 qj = q0;
        const int tj = ONETYPE ? 1 : xq_type(qj);
        double gx = 0, gy = 0, gz = 0;
#pragma unroll
        for (int k = 0; k < CL; k++) {
          const double dx = xi[k] - qj.x, dy = yi[k] - qj.y, dz = zi[k] - qj.z;
          // list membership is exact in the mask (k_build_cluster tests rsq without FMA like the reference); the force
          // cutoff test and the force itself use the fused form (the FP64 pipe is this kernel's top unit)
          const double rsq = fma(dz, dz, fma(dy, dy, dx * dx));
          double cutsq, lj1, lj2, lj3, lj4, offset;
          if (ONETYPE) { cutsq = one.cutsq; lj1 = one.lj1; lj2 = one.lj2; lj3 = one.lj3; lj4 = one.lj4; offset = one.offset; }
          else {
            const double *t = &s_tab[(ti[k] * nt1 + tj) * 6];
            cutsq = t[0]; lj1 = t[1]; lj2 = t[2]; lj3 = t[3]; lj4 = t[4]; offset = t[5];
          }
          // branch-free: the four pair evaluations of a slot interleave; a masked-out pair is evaluated at rsq = 1
          // and multiplied by zero
          const bool hit = ((m >> k) & 1u) && rsq < cutsq;
          const double r2inv = rcp_nr(hit ? rsq : 1.0);
          const double r6inv = r2inv * r2inv * r2inv;
          const double forcelj = r6inv * (lj1 * r6inv - lj2);
          const double fpair = hit ? forcelj * r2inv : 0.0;
          fx[k] = fma(dx, fpair, fx[k]); fy[k] = fma(dy, fpair, fy[k]); fz[k] = fma(dz, fpair, fz[k]);
          if (HALF) { gx = fma(dx, fpair, gx); gy = fma(dy, fpair, gy); gz = fma(dz, fpair, gz); }
          if (EV) {
            const double px = dx * fpair, py = dy * fpair, pz = dz * fpair;
            ev[0] += hit ? r6inv * (lj3 * r6inv - lj4) - offset : 0.0;
            ev[1] += dx * px; ev[2] += dy * py; ev[3] += dz * pz;
            ev[4] += dx * py; ev[5] += dx * pz; ev[6] += dy * pz;
          }
        }
        if (HALF && m && (gx != 0.0 || gy != 0.0 || gz != 0.0)) {
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
    // lane k of the group stores atom k (L >= 4), or lane 0 all of them
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

template <int L, bool EV, bool HALF>
static void launch_lj_cl2(b200md_ctx *c, int nblocks, const LJOne &one, const int *tiles, int ntl) {
  const int ncl = list_units(c);
  if (c->ntypes == 1)
    k_lj_cl<L, EV, HALF, true><<<nblocks, LJ_CL_THREADS, 0, c->stream>>>(c->nlocal, ncl, c->nmax, c->xq, c->neigh, c->numneigh,
                                                                      c->pitch, one, c->ljtab, c->ntypes, c->f, c->partials, tiles, ntl);
  else
    k_lj_cl<L, EV, HALF, false><<<nblocks, LJ_CL_THREADS, 0, c->stream>>>(c->nlocal, ncl, c->nmax, c->xq, c->neigh, c->numneigh,
                                                                       c->pitch, one, c->ljtab, c->ntypes, c->f, c->partials, tiles, ntl);
}
template <int L>
static void launch_lj_cl(b200md_ctx *c, int nblocks, bool ev, const LJOne &one, const int *tiles, int ntl) {
  const bool half = c->list_kind == B200MD_LIST_HALF_NEWTON;
  if (ev) { if (half) launch_lj_cl2<L, true, true>(c, nblocks, one, tiles, ntl); else launch_lj_cl2<L, true, false>(c, nblocks, one, tiles, ntl); }
  else { if (half) launch_lj_cl2<L, false, true>(c, nblocks, one, tiles, ntl); else launch_lj_cl2<L, false, false>(c, nblocks, one, tiles, ntl); }
}

// one launch over all tiles (tiles == nullptr) or over a tile list (interior / boundary split)
int pair_lj_launch(b200md_ctx *c, bool ev, const int *tiles, int ntl) {
  const int ntiles = tiles ? ntl : list_tiles(c);
  const int tpb = (c->cluster ? LJ_CL_THREADS : PAIR_THREADS) / 32;   // tiles (warps) per block
  const int nblocks = (ntiles + tpb - 1) / tpb;
  if (nblocks == 0) return B200MD_OK;
  LJOne one = {0, 0, 0, 0, 0, 0};
  if (c->ntypes == 1) {
    const double *t = &c->h_ljtab[((size_t)1 * 2 + 1) * 8];  // row (1,1)
    one = {t[0], t[1], t[2], t[3], t[4], t[5]};
  }
  if (c->cluster) switch (c->tpa) {
    case 1: launch_lj_cl<1>(c, nblocks, ev, one, tiles, ntl); break;
    case 2: launch_lj_cl<2>(c, nblocks, ev, one, tiles, ntl); break;
    case 4: launch_lj_cl<4>(c, nblocks, ev, one, tiles, ntl); break;
    case 8: launch_lj_cl<8>(c, nblocks, ev, one, tiles, ntl); break;
    case 16: launch_lj_cl<16>(c, nblocks, ev, one, tiles, ntl); break;
    case 32: launch_lj_cl<32>(c, nblocks, ev, one, tiles, ntl); break;
    default: b200md_set_error("b200md: lanes per atom must be 1, 2, 4, 8, 16 or 32"); return B200MD_ERR;
  }
  else switch (c->tpa) {
    case 1: launch_lj<1>(c, nblocks, ev, one, tiles, ntl); break;
    case 2: launch_lj<2>(c, nblocks, ev, one, tiles, ntl); break;
    case 4: launch_lj<4>(c, nblocks, ev, one, tiles, ntl); break;
    case 8: launch_lj<8>(c, nblocks, ev, one, tiles, ntl); break;
    case 16: launch_lj<16>(c, nblocks, ev, one, tiles, ntl); break;
    case 32: launch_lj<32>(c, nblocks, ev, one, tiles, ntl); break;
    default: b200md_set_error("b200md: lanes per atom must be 1, 2, 4, 8, 16 or 32"); return B200MD_ERR;
  }
  c->launches++;
  c->pair_launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

int pair_lj_compute(b200md_ctx *c, int eflag, int vflag) {
  if (c->ntypes > LJ_MAX_TYPES_SMEM) { b200md_set_error("b200md: lj/cut supports at most 8 atom types"); return B200MD_ERR; }
  const bool ev = eflag || vflag;
  const int ntiles = list_tiles(c);
  const int tpb = (c->cluster ? LJ_CL_THREADS : PAIR_THREADS) / 32;
  const int nblocks = (ntiles + tpb - 1) / tpb;
  const bool half = c->list_kind == B200MD_LIST_HALF_NEWTON;
  if (half && c->nlocal + c->nghost > 0)  // Verlet::force_clear over owned + ghost (src/verlet.cpp:376-421): 3 rows, one launch
    CUDA_TRY(cudaMemset2DAsync(c->f, sizeof(double) * (size_t)c->nmax, 0, sizeof(double) * (c->nlocal + c->nghost), 3, c->stream));
  c->ev_valid = false;
  if (nblocks == 0) {
    if (ev) { c->eng_vdwl = 0; for (int k = 0; k < 6; k++) c->virial[k] = 0; c->ev_valid = true; }
    return B200MD_OK;
  }
  if (ev) TRY(ensure_partials(c, nblocks));
  TRY(pair_lj_launch(c, ev, nullptr, 0));
  if (ev) {
    double out[7];
    TRY(reduce_partials(c, nblocks, 7, out));
    const double w = half ? 1.0 : 0.5;
    c->eng_vdwl = w * out[0];
    for (int k = 0; k < 6; k++) c->virial[k] = w * out[1 + k];
    c->ev_valid = true;
  }
  return B200MD_OK;
}
