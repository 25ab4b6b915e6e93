// nve.cu — fix nve integration, rebuild-distance check and thermo reductions.
//
// Reference (/root/reference/src):
//   FixNVE::initial_integrate  fix_nve.cpp:64-104   dtfm = dtf/mass[type]; v += dtfm*f; x += dtv*v
//   FixNVE::final_integrate    fix_nve.cpp:108-141  v += dtfm*f
//   Neighbor::check_distance   neighbor.cpp:2338-2390  any |x-xhold|^2 > triggersq (fused into
//                              the initial-integrate stream: x is in registers already)
//   ComputeTemp::compute_scalar compute_temp.cpp:72-99  sum mass*v.v
// The two half-kicks are kept separate (not the Kokkos fused 2*dtf kick,
// src/KOKKOS/fix_nve_kokkos.cpp:195-205) and use unfused mul/add so trajectories round like
// the FMA-free CPU reference.  These are pure streams: 152 B (initial, +24 with the check)
// and 72 B (final) per atom.
#include "ctx.h"
#include "util.cuh"

template <bool CHECK>
__global__ void k_nve_initial(int n, int nmax, double4 *__restrict__ xq, double *__restrict__ v,
                              const double *__restrict__ f, const double *__restrict__ mass,
                              const double *__restrict__ xhold, double dtv, double dtf, double triggersq,
                              int *__restrict__ flags) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 q = xq[i];
  const double dtfm = dtf / mass[xq_type(q)];
  const size_t s = (size_t)nmax;
  double vx = v[i], vy = v[i + s], vz = v[i + 2 * s];
  vx = __dadd_rn(vx, __dmul_rn(dtfm, f[i]));
  vy = __dadd_rn(vy, __dmul_rn(dtfm, f[i + s]));
  vz = __dadd_rn(vz, __dmul_rn(dtfm, f[i + 2 * s]));
  q.x = __dadd_rn(q.x, __dmul_rn(dtv, vx));
  q.y = __dadd_rn(q.y, __dmul_rn(dtv, vy));
  q.z = __dadd_rn(q.z, __dmul_rn(dtv, vz));
  v[i] = vx; v[i + s] = vy; v[i + 2 * s] = vz;
  xq[i] = q;
  if (CHECK) {
    const double dx = q.x - xhold[i], dy = q.y - xhold[i + s], dz = q.z - xhold[i + 2 * s];
    if (rsq_nofma(dx, dy, dz) > triggersq) flags[0] = 1;
  }
}

__global__ void k_nve_final(int n, int nmax, const double4 *__restrict__ xq, double *__restrict__ v,
                            const double *__restrict__ f, const double *__restrict__ mass, double dtf) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double dtfm = dtf / mass[xq_type(xq[i])];
  const size_t s = (size_t)nmax;
  v[i] = __dadd_rn(v[i], __dmul_rn(dtfm, f[i]));
  v[i + s] = __dadd_rn(v[i + s], __dmul_rn(dtfm, f[i + s]));
  v[i + 2 * s] = __dadd_rn(v[i + 2 * s], __dmul_rn(dtfm, f[i + 2 * s]));
}

// FixNVE::final_integrate of step n and FixNVE::initial_integrate of step n+1 in one pass
// (opt-in, B200MD_FUSE_NVE=1).  Both use the same forces, so the pair is
//   v = v + dtfm f;  v = v + dtfm f;  x = x + dtv v
// with every rounding kept where the two separate kernels have it: results are bit-identical,
// the v / f / type streams are read once instead of twice (72 of 192 B per atom-step saved).
// Only between steps with nothing in between: no thermo (needs v after the first kick), no
// thermostat.
template <bool CHECK>
__global__ void k_nve_final_initial(int n, int nmax, double4 *__restrict__ xq, double *__restrict__ v,
                                    const double *__restrict__ f, const double *__restrict__ mass,
                                    const double *__restrict__ xhold, double dtv, double dtf, double triggersq,
                                    int *__restrict__ flags) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 q = xq[i];
  const double dtfm = dtf / mass[xq_type(q)];
  const size_t s = (size_t)nmax;
  const double kx = __dmul_rn(dtfm, f[i]), ky = __dmul_rn(dtfm, f[i + s]), kz = __dmul_rn(dtfm, f[i + 2 * s]);
  double vx = __dadd_rn(__dadd_rn(v[i], kx), kx);
  double vy = __dadd_rn(__dadd_rn(v[i + s], ky), ky);
  double vz = __dadd_rn(__dadd_rn(v[i + 2 * s], kz), kz);
  q.x = __dadd_rn(q.x, __dmul_rn(dtv, vx));
  q.y = __dadd_rn(q.y, __dmul_rn(dtv, vy));
  q.z = __dadd_rn(q.z, __dmul_rn(dtv, vz));
  v[i] = vx; v[i + s] = vy; v[i + 2 * s] = vz;
  xq[i] = q;
  if (CHECK) {
    const double dx = q.x - xhold[i], dy = q.y - xhold[i + s], dz = q.z - xhold[i + 2 * s];
    if (rsq_nofma(dx, dy, dz) > triggersq) flags[0] = 1;
  }
}

int nve_final_initial(b200md_ctx *c, double dtv, double dtf, bool check_dist) {
  const int n = c->nlocal;
  if (n == 0) return B200MD_OK;
  if (check_dist)
    k_nve_final_initial<true><<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->nmax, c->xq, c->v, c->f, c->mass, c->xhold, dtv,
                                                                     dtf, c->triggersq, c->d_flags);
  else
    k_nve_final_initial<false><<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->nmax, c->xq, c->v, c->f, c->mass, c->xhold, dtv,
                                                                      dtf, c->triggersq, c->d_flags);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

int nve_initial(b200md_ctx *c, double dtv, double dtf, bool check_dist) {
  const int n = c->nlocal;
  if (n == 0) return B200MD_OK;
  if (check_dist)
    k_nve_initial<true><<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->nmax, c->xq, c->v, c->f, c->mass, c->xhold, dtv,
                                                               dtf, c->triggersq, c->d_flags);
  else
    k_nve_initial<false><<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->nmax, c->xq, c->v, c->f, c->mass, c->xhold, dtv,
                                                                dtf, c->triggersq, c->d_flags);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

int nve_final(b200md_ctx *c, double dtf) {
  const int n = c->nlocal;
  if (n == 0) return B200MD_OK;
  k_nve_final<<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->nmax, c->xq, c->v, c->f, c->mass, dtf);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

// ---- reductions ----
int ensure_partials(b200md_ctx *c, int nblocks) {
  if (nblocks <= c->partial_cap) return B200MD_OK;
  if (c->partials) CUDA_TRY(cudaFree(c->partials));
  c->partial_cap = nblocks + nblocks / 4 + 64;
  CUDA_TRY(cudaMalloc(&c->partials, sizeof(double) * 8 * (size_t)c->partial_cap));
  return B200MD_OK;
}

// one block sums nblocks rows of NV values in a fixed order (deterministic)
__global__ void __launch_bounds__(1024) k_reduce_rows(const double *__restrict__ partials, int nblocks, int nvals,
                                                      double *__restrict__ out) {
  __shared__ double s[32];
  for (int k = 0; k < nvals; k++) {
    double t = 0;
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) t += partials[(size_t)b * nvals + k];
    t = group_sum<32>(t);
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x < 32) {
      double u = threadIdx.x < (blockDim.x >> 5) ? s[threadIdx.x] : 0.0;
      u = group_sum<32>(u);
      if (threadIdx.x == 0) out[k] = u;
    }
    __syncthreads();
  }
}

int reduce_partials(b200md_ctx *c, int nblocks, int nvals, double *host_out) {
  k_reduce_rows<<<1, 1024, 0, c->stream>>>(c->partials, nblocks, nvals, c->d_result);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaMemcpyAsync(c->h_result, c->d_result, sizeof(double) * nvals, cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  for (int k = 0; k < nvals; k++) host_out[k] = c->h_result[k];
  return B200MD_OK;
}

// sums per block: m*v.v, m*vx, m*vy, m*vz, m  (temperature and momentum zeroing)
__global__ void __launch_bounds__(256) k_ke(int n, int nmax, const double4 *__restrict__ xq, const double *__restrict__ v,
                                            const double *__restrict__ mass, double *__restrict__ partials) {
  double val[5] = {0, 0, 0, 0, 0};
  const size_t s = (size_t)nmax;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double m = mass[xq_type(xq[i])];
    const double vx = v[i], vy = v[i + s], vz = v[i + 2 * s];
    val[0] += (vx * vx + vy * vy + vz * vz) * m;  // compute_temp.cpp:88-90
    val[1] += m * vx; val[2] += m * vy; val[3] += m * vz; val[4] += m;
  }
  block_reduce_store<5>(val, partials);
}

int thermo_ke(b200md_ctx *c, double *sums5) {
  const int n = c->nlocal;
  int nblocks = std::min((n + 255) / 256, B200_SMS * 8);
  if (nblocks < 1) nblocks = 1;
  TRY(ensure_partials(c, nblocks));
  k_ke<<<nblocks, 256, 0, c->stream>>>(n, c->nmax, c->xq, c->v, c->mass, c->partials);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  return reduce_partials(c, nblocks, 5, sums5);
}

__global__ void k_vel_adjust(int n, int nmax, double *__restrict__ v, double sx, double sy, double sz, double factor) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const size_t s = (size_t)nmax;
  v[i] = (v[i] - sx) * factor; v[i + s] = (v[i + s] - sy) * factor; v[i + 2 * s] = (v[i + 2 * s] - sz) * factor;
}
int vel_adjust(b200md_ctx *c, double sx, double sy, double sz, double factor) {
  if (c->nlocal == 0) return B200MD_OK;
  k_vel_adjust<<<(c->nlocal + 255) / 256, 256, 0, c->stream>>>(c->nlocal, c->nmax, c->v, sx, sy, sz, factor);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}
