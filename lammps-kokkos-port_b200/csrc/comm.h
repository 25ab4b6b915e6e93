// comm.h — NCCL state and helpers shared by comm.cu (staged halo, single-rank fast path) and halo26.cu (direct halo)
#pragma once
#include <nccl.h>
#include <algorithm>
#include "ctx.h"
#include "util.cuh"

struct NcclState {
  ncclComm_t comm = nullptr;
  int *d_counts = nullptr;  // [4] scratch for count exchange
  int *h_counts = nullptr;  // pinned [4]
  double *d_red = nullptr;  // [16] allreduce scratch
};

#define NCCL_TRY(call)                                                                    \
  do {                                                                                    \
    ncclResult_t _r = (call);                                                             \
    if (_r != ncclSuccess) {                                                              \
      b200md_set_error(std::string("b200md: NCCL error: ") + ncclGetErrorString(_r) +     \
                       " at " + __FILE__ + ":" + std::to_string(__LINE__));               \
      return B200MD_ERR;                                                                  \
    }                                                                                     \
  } while (0)


// x + pbc*prd once per crossed image like AtomVec::pack_comm/pack_border (atom_vec.cpp:349-378, 794-830)
__device__ __forceinline__ double shift_n(double x, int n, double prd) {
  for (; n > 0; n--) x = __dadd_rn(x, prd);
  for (; n < 0; n++) x = __dsub_rn(x, prd);
  return x;
}

static inline int ensure_buf(double *&p, size_t &cap, size_t need_doubles) {
  if (need_doubles <= cap) return B200MD_OK;
  if (p) CUDA_TRY(cudaFree(p));
  cap = need_doubles + need_doubles / 4 + 1024;
  CUDA_TRY(cudaMalloc(&p, cap * sizeof(double)));
  return B200MD_OK;
}
static inline int nblk(int n) { return (n + 255) / 256; }

// halo26.cu
int h26_setup(b200md_ctx *c, bool want);
int h26_exchange(b200md_ctx *c);
int h26_borders(b200md_ctx *c);
int h26_forward(b200md_ctx *c, double *scalar);   // scalar == nullptr: positions
int h26_reverse3(b200md_ctx *c, double *arr);
int h26_reverse1(b200md_ctx *c, double *arr);
void h26_free(b200md_ctx *c);
