// util.cuh — device helpers shared by the kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define B200_SMS 148

__device__ __forceinline__ int xq_type(const double4 &q) {
  return (int)(__double_as_longlong(q.w) & 0xffffffffll);
}
__device__ __forceinline__ double type_bits(int t) { return __longlong_as_double((long long)t); }

// rsq in the reference's operand order, no FMA contraction:
// delx*delx + dely*dely + delz*delz  (src/npair_full_bin_atomonly.cpp:74-77)
__device__ __forceinline__ double rsq_nofma(double dx, double dy, double dz) {
  return __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz));
}

// one 32-byte sector gather of {x,y,z,type}
__device__ __forceinline__ double4 ld_xq(const double4 *p) {
  double4 r;
#if defined(B200MD_LD256)
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];"
               : "=d"(r.x), "=d"(r.y), "=d"(r.z), "=d"(r.w)
               : "l"(p));
#else
  const double2 *q = reinterpret_cast<const double2 *>(p);
  double2 a = __ldg(q), b = __ldg(q + 1);
  r.x = a.x; r.y = a.y; r.z = b.x; r.w = b.y;
#endif
  return r;
}

template <int W>
__device__ __forceinline__ double group_sum(double v) {
#pragma unroll
  for (int o = W / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// block reduction of NV doubles per thread into out[blockIdx.x*NV + k]; blockDim multiple of 32
template <int NV>
__device__ __forceinline__ void block_reduce_store(double (&val)[NV], double *out) {
  __shared__ double s_red[32][NV];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int nw = (blockDim.x + 31) >> 5;
#pragma unroll
  for (int k = 0; k < NV; k++) val[k] = group_sum<32>(val[k]);
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) s_red[wid][k] = val[k];
  }
  __syncthreads();
  if (wid == 0) {
#pragma unroll
    for (int k = 0; k < NV; k++) {
      double t = lane < nw ? s_red[lane][k] : 0.0;
      t = group_sum<32>(t);
      if (lane == 0) out[(size_t)blockIdx.x * NV + k] = t;
    }
  }
}

// cluster-pair list (neigh.cu): CL owned atoms share one j-list, entry = j | mask << CL_SHIFT
#ifndef CL
#define CL 4
#endif
#define CL_FULL ((1u << CL) - 1u)
#define CL_SHIFT 28
#define CL_JMASK 0x0fffffff

// 1/x to FP64 rounding error (not correctly rounded): MUFU.RCP64H seed (>= 20 bits) + ONE third-order step,
// y (1 + e + e^2) with e = 1 - x y: the error goes from e to e^3 (2^-60 or better) in three FMAs, where two Newton
// steps take four.  x is a squared pair distance (normal, positive); the IEEE division is ~12 instructions.
__device__ __forceinline__ double rcp_nr(double x) {
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double e = fma(-x, y, 1.0);
  const double t = fma(e, e, e);
  return fma(y, t, y);
}

// 1/sqrt(x) to FP64 rounding error (not correctly rounded): MUFU.RSQ64H seed + two Newton steps
__device__ __forceinline__ double rsqrt_nr(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double xh = 0.5 * x;
  double e = fma(-xh * y, y, 0.5);
  y = fma(y, e, y);
  e = fma(-xh * y, y, 0.5);
  return fma(y, e, y);
}

// 32-byte table row fetched only by the lanes whose predicate holds (others keep zeros and issue no L1 wavefront):
// the cluster kernels evaluate masked-out slots branch-free, but must not pay table look-ups for them
__device__ __forceinline__ double4 ld4_if(const double *p, bool on) {
  double4 r = make_double4(0.0, 0.0, 0.0, 0.0);
  asm volatile("{\n\t.reg .pred q;\n\tsetp.ne.u32 q, %5, 0;\n\t@q ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];\n\t}"
               : "+d"(r.x), "+d"(r.y), "+d"(r.z), "+d"(r.w)
               : "l"(p), "r"((unsigned)on));
  return r;
}
