// ke_static.cu — b200md_ke_sums_static: the kinetic sums of thermo (k_ke + k_reduce_rows, nve.cu) for callers that
// must not allocate.  No counterpart in the reference (include/b200md.h).
//
// The existing path (thermo_ke, nve.cu) writes one row of partial sums per block into c->partials, which
// ensure_partials() grows with cudaMalloc, and reduces the rows in a second launch.  Here the rows live in a
// file-scope static device array of fixed size, and the last block to finish reduces them (one launch, no host
// synchronisation, nothing allocated): usable under stream capture, in a serving loop, or by a harness that
// forbids cudaMalloc.  Every GPU that loads the library holds the array (KE_STATIC_BYTES) until exit.
//
// The array is shared by every call on a device and keeps the last call's contents.  What must start at zero is the
// arrival counter: the reducing block resets it before it exits, and the set-up clears it once.  Two calls that
// overlap on the GPU (different streams, no ordering between them) race on the rows and on the counter: order them.
#include "ctx.h"
#include "util.cuh"

#include <mutex>

#define KE_STATIC_THREADS 256
#define KE_STATIC_ATOMS_PER_BLOCK 4096                                        // 16 atoms per thread
#define KE_STATIC_MAX_BLOCKS (B200MD_KE_STATIC_MAX_ATOMS / KE_STATIC_ATOMS_PER_BLOCK)   // 4096
#define KE_STATIC_NV 5
#define KE_STATIC_DOUBLES (KE_STATIC_MAX_BLOCKS * KE_STATIC_NV + 32)          // rows + one 256-byte line for the counter
#define KE_STATIC_BYTES (KE_STATIC_DOUBLES * 8)                               // 164 096 bytes

static __device__ __align__(256) double g_ke_scratch[KE_STATIC_DOUBLES];

// In host code g_ke_scratch is only a handle for cudaGetSymbolAddress: the device address is looked up once per
// device (with that device current) and handed to the kernel as an argument.
#define KE_STATIC_MAX_DEVICES 64
static double *g_ke_ptr[KE_STATIC_MAX_DEVICES];
static std::mutex g_ke_mutex;

static int ke_static_setup(int device, double **ptr) {
  std::lock_guard<std::mutex> lock(g_ke_mutex);
  if (!g_ke_ptr[device]) {
    double *p = nullptr;
    CUDA_TRY(cudaGetSymbolAddress((void **)&p, g_ke_scratch));
    if (!p) { b200md_set_error("b200md_ke_sums_static: no device address for the scratch array"); return B200MD_ERR; }
    // the arrival counter (last 256-byte line); a plain synchronous memset, once per device
    CUDA_TRY(cudaMemset(p + KE_STATIC_MAX_BLOCKS * KE_STATIC_NV, 0, 256));
    g_ke_ptr[device] = p;
  }
  *ptr = g_ke_ptr[device];
  return B200MD_OK;
}

// Block b sums atoms [b*4096, (b+1)*4096): m*v.v, m*vx, m*vy, m*vz, m (compute_temp.cpp:88-90 and the momentum sums
// of velocity.cpp).  The block -> atoms map and the order of the final reduction depend on n only, so the result
// is the same bits on every call with the same input.
__global__ void __launch_bounds__(KE_STATIC_THREADS)
k_ke_static(int n, size_t stride, const double *__restrict__ v, const int *__restrict__ type,
            const double *__restrict__ mass, double *__restrict__ scratch, double *__restrict__ out5) {
  __shared__ bool s_last;
  double val[KE_STATIC_NV] = {0, 0, 0, 0, 0};
  const int lo = blockIdx.x * KE_STATIC_ATOMS_PER_BLOCK;
  const int hi = min(n, lo + KE_STATIC_ATOMS_PER_BLOCK);
  for (int i = lo + threadIdx.x; i < hi; i += KE_STATIC_THREADS) {
    const double m = mass[type ? type[i] : 1];
    const double vx = v[i], vy = v[i + stride], vz = v[i + 2 * stride];
    val[0] += (vx * vx + vy * vy + vz * vz) * m;
    val[1] += m * vx; val[2] += m * vy; val[3] += m * vz; val[4] += m;
  }
  block_reduce_store<KE_STATIC_NV>(val, scratch);   // row blockIdx.x
  unsigned int *counter = reinterpret_cast<unsigned int *>(scratch + KE_STATIC_MAX_BLOCKS * KE_STATIC_NV);
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();                                 // the row is visible before the arrival is
    s_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // last block: rows in a fixed order (thread t takes rows t, t+256, ...), then the block tree
  double tot[KE_STATIC_NV] = {0, 0, 0, 0, 0};
  for (int r = threadIdx.x; r < (int)gridDim.x; r += KE_STATIC_THREADS) {
#pragma unroll
    for (int k = 0; k < KE_STATIC_NV; k++) tot[k] += __ldcg(scratch + (size_t)r * KE_STATIC_NV + k);
  }
  __shared__ double s_tot[KE_STATIC_THREADS / 32][KE_STATIC_NV];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
  for (int k = 0; k < KE_STATIC_NV; k++) tot[k] = group_sum<32>(tot[k]);
  if (lane == 0) {
#pragma unroll
    for (int k = 0; k < KE_STATIC_NV; k++) s_tot[wid][k] = tot[k];
  }
  __syncthreads();
  if (threadIdx.x < KE_STATIC_NV) {
    double t = 0.0;
    for (int w = 0; w < KE_STATIC_THREADS / 32; w++) t += s_tot[w][threadIdx.x];
    out5[threadIdx.x] = t;
  }
  if (threadIdx.x == 0) *counter = 0u;               // the next call starts from zero arrivals
}

extern "C" int b200md_ke_sums_static(int n, long long stride, const double *v_soa_dev, const int *type_dev,
                                     const double *mass_dev, double *out5_dev, void *stream) {
  if (n < 0 || n > B200MD_KE_STATIC_MAX_ATOMS) {
    b200md_set_error("b200md_ke_sums_static: n = " + std::to_string(n) + " is outside 0.." +
                     std::to_string(B200MD_KE_STATIC_MAX_ATOMS) +
                     " (B200MD_KE_STATIC_MAX_ATOMS, the size of the static scratch array)");
    return B200MD_ERR;
  }
  if (stride < n) { b200md_set_error("b200md_ke_sums_static: stride is smaller than n"); return B200MD_ERR; }
  int device = 0;
  CUDA_TRY(cudaGetDevice(&device));
  if (device < 0 || device >= KE_STATIC_MAX_DEVICES) {
    b200md_set_error("b200md_ke_sums_static: device index beyond the per-device address table");
    return B200MD_ERR;
  }
  double *scratch = nullptr;
  TRY(ke_static_setup(device, &scratch));
  if (n == 0) return B200MD_OK;                      // set-up only (call once outside a stream capture)
  if (!v_soa_dev || !mass_dev || !out5_dev) { b200md_set_error("b200md_ke_sums_static: null pointer"); return B200MD_ERR; }
  const int nblocks = (n + KE_STATIC_ATOMS_PER_BLOCK - 1) / KE_STATIC_ATOMS_PER_BLOCK;
  k_ke_static<<<nblocks, KE_STATIC_THREADS, 0, (cudaStream_t)stream>>>(n, (size_t)stride, v_soa_dev, type_dev, mass_dev,
                                                                      scratch, out5_dev);
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}
