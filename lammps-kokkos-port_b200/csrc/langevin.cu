// langevin.cu — fix langevin on the device, with the reference's exact RanMars stream.
//
// Reference (/root/reference/src):
//   FixLangevin::FixLangevin        fix_langevin.cpp:60-150   RanMars(seed + comm->me), ratio[]
//   FixLangevin::init               fix_langevin.cpp:286-298  gfactor1 / gfactor2 per type
//   FixLangevin::setup              fix_langevin.cpp:360-361  post_force on the set-up forces
//   FixLangevin::compute_target     fix_langevin.cpp:773-787  T ramp over the run window
//   post_force_templated<0,..,0>    fix_langevin.cpp:576-713  f += gamma1 v + gamma2 (u - 0.5),
//                                   three draws per atom, atoms in local index order
//   RanMars                         random_mars.cpp:30-95     lagged Fibonacci (97, 33) minus an
//                                   arithmetic sequence, all in multiples of 2^-24
//
// B200 design.  The reference's thermostat is serial by construction: one generator, drawn
// atom after atom.  Its recurrence is linear over Z/2^24, though (host.cpp, "RanMars in integer
// form"), so the state k draws ahead is a fixed 97x97 integer matrix times the current state.
// Each force evaluation needs the next S = 3*nlocal draws, cut into chunks of CHUNK.  Chunk m of
// every evaluation is served by block m, which keeps its OWN generator state resident in HBM:
//   set-up (once, k_ranmars_chunk_states): state_m = J^(m*CHUNK) state, by binary decomposition
//               of m over precomputed J^(CHUNK*2^b) tables (<= 24 mat-vecs of 97x97)
//   every evaluation (k_ranmars_fill): block m
//            -> CHUNK values in shared memory, 32 per warp iteration (lag 33 > 32, so a warp's
//               lanes never depend on each other inside an iteration)
//            -> subtracts c[n] in closed form and stores 24-bit integers
//            -> state_m = J^S state_m: ONE mat-vec with the one matrix every block shares
//               (37 KB, L1/L2-resident), because the same chunk of the next evaluation starts
//               exactly S draws later — no serial dependency between blocks, ever.
// The thermostat kernel then reads its three numbers at the atom's HOST rank (the local index
// the reference would have drawn them at; rank_by_tag is built from the last b200md_set_atoms
// order), so the device's own bin-sorted atom order does not matter.  In the step loop it is
// fused with FixNVE::final_integrate: one pass over v and f instead of two.
// Per atom-step: 12 B of random words written and read, 4 B tag, 4 B rank, plus the 96 B of the
// v / f streams nve_final moves anyway.
#include <cub/device/device_radix_sort.cuh>
#include "ctx.h"
#include "util.cuh"

#define RM_LAG B200MD_RANMARS_LAG
#define RM_CHUNK 2048
#define RM_LEVELS 24
#define RM_THREADS 128
#define RM_MASK 0xffffffu

void ranmars_jump_matrix(long long k, uint32_t *M);   // host.cpp

// c[n] of draw n (n = 0 is the constructor's own draw), in units of 2^-24
__device__ __forceinline__ uint32_t mars_c(long long n) {
  const long long CM = 16777213LL;
  const long long t = (((n + 1) % CM) * 7654321LL) % CM;
  return (uint32_t)((362436LL + CM - t) % CM);
}

// thread `tid` < 97 computes row tid of (matrix^T given transposed) x vector; MT[j*97 + tid] so
// consecutive threads read consecutive words
__device__ __forceinline__ uint32_t matvec_row(const uint32_t *__restrict__ MT, const uint32_t *v, int tid) {
  uint32_t acc = 0;
#pragma unroll 16
  for (int j = 0; j < RM_LAG; j++) acc += __ldg(MT + j * RM_LAG + tid) * v[j];
  return acc & RM_MASK;
}

// jumpT: [RM_LEVELS][97][97], level b = TRANSPOSE of the jump-by-(CHUNK*2^b) matrix.
// chunk_state[m] = J^(m*CHUNK) state_in
__global__ void __launch_bounds__(RM_THREADS)
k_ranmars_chunk_states(const uint32_t *__restrict__ state_in, const uint32_t *__restrict__ jumpT,
                       uint32_t *__restrict__ chunk_state) {
  __shared__ uint32_t st[2][RM_LAG];
  const int tid = threadIdx.x;
  const int m = blockIdx.x;
  int cur = 0;
  if (tid < RM_LAG) st[0][tid] = state_in[tid];
  __syncthreads();
  for (int b = 0; b < RM_LEVELS; b++) {
    if (!((m >> b) & 1)) continue;   // uniform across the block
    if (tid < RM_LAG) st[cur ^ 1][tid] = matvec_row(jumpT + (size_t)b * RM_LAG * RM_LAG, st[cur], tid);
    __syncthreads();
    cur ^= 1;
  }
  if (tid < RM_LAG) chunk_state[(size_t)m * RM_LAG + tid] = st[cur][tid];
}

// draws [position + m*CHUNK, ...) of this evaluation from block m's resident state, then
// state_m <- J^S state_m (stepT = transposed jump-by-S matrix, S = draws per evaluation)
__global__ void __launch_bounds__(RM_THREADS)
k_ranmars_fill(uint32_t *__restrict__ chunk_state, const uint32_t *__restrict__ stepT, long long position, int nnum,
               uint32_t *__restrict__ out) {
  __shared__ uint32_t buf[RM_LAG + RM_CHUNK];
  const int tid = threadIdx.x;
  const int m = blockIdx.x;
  uint32_t *mine = chunk_state + (size_t)m * RM_LAG;
  if (tid < RM_LAG) buf[tid] = mine[tid];
  __syncthreads();
  // the state hop does not depend on the generation: issue its loads first
  uint32_t next = 0;
  if (tid < RM_LAG) next = matvec_row(stepT, buf, tid);
  const int base = m * RM_CHUNK;                 // first draw of this chunk, relative to `position`
  const int want = min(RM_CHUNK, nnum - base);   // > 0 by the grid size
  if (tid < 32) {
    const int iters = (want + 31) / 32;
    for (int t = 0; t < iters; t++) {
      const int k = 32 * t + tid;
      buf[RM_LAG + k] = (buf[k] - buf[k + 64]) & RM_MASK;   // L[n+97] = L[n] - L[n+64]
      __syncwarp();
    }
  }
  __syncthreads();
  for (int k = tid; k < want; k += RM_THREADS)
    out[base + k] = (buf[RM_LAG + k] - mars_c(position + base + k)) & RM_MASK;
  if (tid < RM_LAG) mine[tid] = next;
}

// f += gamma1 v + gamma2 (u - 0.5)  [+ FixNVE::final_integrate when FINAL]
// NEXT (with FINAL): FixNVE::initial_integrate of the following step in the same pass, like k_nve_final_initial (nve.cu):
// both half-kicks use the same thermostatted force, every rounding stays where the separate kernels have it, and the
// rebuild-distance check of the next step rides along (CHECK).  Only between steps with nothing in between.
template <bool FINAL, bool NEXT, bool CHECK>
__global__ void k_langevin(int n, int nmax, double4 *__restrict__ xq, double *__restrict__ v,
                           double *__restrict__ f, const int *__restrict__ tag, const int *__restrict__ rank_by_tag,
                           const uint32_t *__restrict__ rnd, const double *__restrict__ gfac /* [ntypes+1][2] */,
                           const double *__restrict__ mass, double tsqrt, double dtf, double dtv,
                           const double *__restrict__ xhold, double triggersq, int *__restrict__ flags) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double4 q = xq[i];
  const int type = xq_type(q);
  const double gamma1 = gfac[2 * type], gamma2 = __dmul_rn(gfac[2 * type + 1], tsqrt);
  const size_t r = (size_t)rank_by_tag[tag[i]] * 3;
  const size_t s = (size_t)nmax;
  double dtfm = 0.0;
  if (FINAL) dtfm = dtf / mass[type];
  double xn[3] = {q.x, q.y, q.z};
#pragma unroll
  for (int d = 0; d < 3; d++) {
    const double u = (double)rnd[r + d] * (1.0 / 16777216.0);
    const double fran = __dmul_rn(gamma2, u - 0.5);
    const double vv = v[i + d * s];
    const double fdrag = __dmul_rn(gamma1, vv);
    const double fn = __dadd_rn(f[i + d * s], __dadd_rn(fdrag, fran));
    f[i + d * s] = fn;
    if (FINAL) {
      const double kick = __dmul_rn(dtfm, fn);
      double vn = __dadd_rn(vv, kick);
      if (NEXT) {
        vn = __dadd_rn(vn, kick);
        xn[d] = __dadd_rn(xn[d], __dmul_rn(dtv, vn));
      }
      v[i + d * s] = vn;
    }
  }
  if (NEXT) {
    q.x = xn[0]; q.y = xn[1]; q.z = xn[2];
    xq[i] = q;
    if (CHECK) {
      const double dx = q.x - xhold[i], dy = q.y - xhold[i + s], dz = q.z - xhold[i + 2 * s];
      if (rsq_nofma(dx, dy, dz) > triggersq) flags[0] = 1;
    }
  }
}

void langevin_free(b200md_ctx *c) {
  Langevin &L = c->lang;
  if (L.state[0]) cudaFree(L.state[0]);
  if (L.state[1]) cudaFree(L.state[1]);
  if (L.jumpT) cudaFree(L.jumpT);
  if (L.stepT) cudaFree(L.stepT);
  if (L.chunk_state) cudaFree(L.chunk_state);
  if (L.rnd) cudaFree(L.rnd);
  if (L.rank_by_tag) cudaFree(L.rank_by_tag);
  if (L.gfac) cudaFree(L.gfac);
  if (L.sort_keys) cudaFree(L.sort_keys);
  if (L.sort_vals) cudaFree(L.sort_vals);
  if (L.sort_tmp) cudaFree(L.sort_tmp);
  L = Langevin();
}

static int ranmars_device_init(b200md_ctx *c, int seed) {
  Langevin &L = c->lang;
  uint32_t st[RM_LAG];
  TRY(b200md_ranmars_state(seed, st));
  if (!L.state[0]) {
    CUDA_TRY(cudaMalloc(&L.state[0], sizeof(uint32_t) * RM_LAG));
    CUDA_TRY(cudaMalloc(&L.state[1], sizeof(uint32_t) * RM_LAG));
  }
  if (!L.jumpT) {
    // J^(CHUNK*2^b), b < LEVELS, transposed; 24 x 37 KB
    std::vector<uint32_t> M((size_t)RM_LAG * RM_LAG), T((size_t)RM_LEVELS * RM_LAG * RM_LAG);
    for (int b = 0; b < RM_LEVELS; b++) {
      ranmars_jump_matrix((long long)RM_CHUNK << b, M.data());
      uint32_t *Tb = T.data() + (size_t)b * RM_LAG * RM_LAG;
      for (int i = 0; i < RM_LAG; i++)
        for (int j = 0; j < RM_LAG; j++) Tb[j * RM_LAG + i] = M[(size_t)i * RM_LAG + j];
    }
    CUDA_TRY(cudaMalloc(&L.jumpT, sizeof(uint32_t) * T.size()));
    CUDA_TRY(cudaMemcpy(L.jumpT, T.data(), sizeof(uint32_t) * T.size(), cudaMemcpyHostToDevice));
  }
  CUDA_TRY(cudaMemcpy(L.state[0], st, sizeof(st), cudaMemcpyHostToDevice));
  memcpy(L.h_state, st, sizeof(st));
  L.h_position = 1;
  L.position = 1;   // the constructor drew once
  L.stride = 0;     // chunk states are laid out at the first fill
  return B200MD_OK;
}

// resident chunk states for evaluations of `nnum` draws each, starting at the current position
static int ranmars_layout(b200md_ctx *c, int nnum) {
  Langevin &L = c->lang;
  const int nblocks = (nnum + RM_CHUNK - 1) / RM_CHUNK;
  if (nblocks >= (1 << RM_LEVELS)) { b200md_set_error("b200md: fix langevin: too many atoms for one random stream"); return B200MD_ERR; }
  // generator state at L.position, from the last state the host knows (set-up time only)
  uint32_t st[RM_LAG];
  TRY(b200md_ranmars_jump(L.h_state, L.position - L.h_position, st));
  memcpy(L.h_state, st, sizeof(st));
  L.h_position = L.position;
  CUDA_TRY(cudaMemcpyAsync(L.state[0], L.h_state, sizeof(st), cudaMemcpyHostToDevice, c->stream));
  if (L.chunk_cap < nblocks) {
    if (L.chunk_state) CUDA_TRY(cudaFree(L.chunk_state));
    L.chunk_cap = nblocks + nblocks / 8 + 8;
    CUDA_TRY(cudaMalloc(&L.chunk_state, sizeof(uint32_t) * RM_LAG * (size_t)L.chunk_cap));
  }
  k_ranmars_chunk_states<<<nblocks, RM_THREADS, 0, c->stream>>>(L.state[0], L.jumpT, L.chunk_state);
  CUDA_TRY(cudaGetLastError());
  c->launches++;
  // the per-evaluation hop J^nnum, transposed
  std::vector<uint32_t> M((size_t)RM_LAG * RM_LAG), T((size_t)RM_LAG * RM_LAG);
  ranmars_jump_matrix(nnum, M.data());
  for (int i = 0; i < RM_LAG; i++)
    for (int j = 0; j < RM_LAG; j++) T[(size_t)j * RM_LAG + i] = M[(size_t)i * RM_LAG + j];
  if (!L.stepT) CUDA_TRY(cudaMalloc(&L.stepT, sizeof(uint32_t) * T.size()));
  CUDA_TRY(cudaStreamSynchronize(c->stream));   // T is a local; also orders against earlier fills
  CUDA_TRY(cudaMemcpy(L.stepT, T.data(), sizeof(uint32_t) * T.size(), cudaMemcpyHostToDevice));
  L.stride = nnum;
  L.nevals = 0;
  return B200MD_OK;
}

// the next nnum draws of the stream into L.rnd (24-bit integers)
static int ranmars_fill(b200md_ctx *c, int nnum) {
  Langevin &L = c->lang;
  if (nnum <= 0) return B200MD_OK;
  if (L.rnd_cap < nnum) {
    if (L.rnd) CUDA_TRY(cudaFree(L.rnd));
    L.rnd_cap = nnum + nnum / 8 + 1024;
    CUDA_TRY(cudaMalloc(&L.rnd, sizeof(uint32_t) * (size_t)L.rnd_cap));
  }
  if (L.stride != nnum) TRY(ranmars_layout(c, nnum));   // first evaluation, or the atom count changed
  const int nblocks = (nnum + RM_CHUNK - 1) / RM_CHUNK;
  k_ranmars_fill<<<nblocks, RM_THREADS, 0, c->stream>>>(L.chunk_state, L.stepT, L.position, nnum, L.rnd);
  CUDA_TRY(cudaGetLastError());
  c->launches++;
  L.position += nnum;
  L.nevals++;
  return B200MD_OK;
}

// rank_by_tag[tag] = index of the atom in the host's order (the order of the last set_atoms)
int langevin_rank_from_host_order(b200md_ctx *c) {
  Langevin &L = c->lang;
  const int n = c->nlocal;
  std::vector<int> tags(n);
  if (n) CUDA_TRY(cudaMemcpy(tags.data(), c->tag, sizeof(int) * n, cudaMemcpyDeviceToHost));
  int maxtag = 0;
  for (int i = 0; i < n; i++) {
    if (tags[i] < 1) { b200md_set_error("b200md: fix langevin requires atom IDs >= 1"); return B200MD_ERR; }
    maxtag = tags[i] > maxtag ? tags[i] : maxtag;
  }
  std::vector<int> rank((size_t)maxtag + 1, -1);
  for (int i = 0; i < n; i++) {
    if (rank[tags[i]] >= 0) { b200md_set_error("b200md: fix langevin requires unique atom IDs"); return B200MD_ERR; }
    rank[tags[i]] = i;
  }
  if (L.rank_by_tag) CUDA_TRY(cudaFree(L.rank_by_tag));
  CUDA_TRY(cudaMalloc(&L.rank_by_tag, sizeof(int) * rank.size()));
  CUDA_TRY(cudaMemcpy(L.rank_by_tag, rank.data(), sizeof(int) * rank.size(), cudaMemcpyHostToDevice));
  L.nrank = n;
  return B200MD_OK;
}

// ---- Atom::sort inside a run (verlet.cpp:285, atom.cpp:2246-2349) -------------------------------
// The reference re-orders its atoms: bins of cutneighmax/2 over the sub-box, stable inside a bin.
// Only the draw each atom gets depends on that order here, so the device re-ranks instead of moving
// anything: key = (bin, previous rank), one radix sort, new rank = position in the sorted sequence.
struct SortBins { double lo[3], binv[3]; int nb[3]; };

__global__ void k_resort_keys(int n, const double4 *__restrict__ xq, const int *__restrict__ tag,
                              const int *__restrict__ rank_by_tag, SortBins G, unsigned long long nrank,
                              unsigned long long *__restrict__ keys, int *__restrict__ vals) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double4 q = xq[i];
  int ix = (int)(__dmul_rn(q.x - G.lo[0], G.binv[0]));
  int iy = (int)(__dmul_rn(q.y - G.lo[1], G.binv[1]));
  int iz = (int)(__dmul_rn(q.z - G.lo[2], G.binv[2]));
  ix = min(max(ix, 0), G.nb[0] - 1); iy = min(max(iy, 0), G.nb[1] - 1); iz = min(max(iz, 0), G.nb[2] - 1);
  const unsigned long long bin = ((unsigned long long)iz * G.nb[1] + iy) * G.nb[0] + ix;
  const int t = tag[i];
  keys[i] = bin * nrank + (unsigned long long)rank_by_tag[t];
  vals[i] = t;
}
__global__ void k_resort_assign(int n, const int *__restrict__ sorted_tag, int *__restrict__ rank_by_tag) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  rank_by_tag[sorted_tag[k]] = k;
}

int langevin_resort(b200md_ctx *c) {
  Langevin &L = c->lang;
  const int n = c->nlocal;
  L.nextsort = (L.now / L.sortfreq) * L.sortfreq + L.sortfreq;   // atom.cpp:2252
  if (n != L.nrank) { b200md_set_error("b200md: fix langevin: the number of atoms changed since it was set"); return B200MD_ERR; }
  // Atom::setup_sort_bins (atom.cpp:2351-2410)
  const double binsize = 0.5 * c->cutneighmax;
  if (binsize <= 0.0 || n == 0) return B200MD_OK;
  SortBins G;
  const double bininv = 1.0 / binsize;
  long long nbins = 1;
  for (int d = 0; d < 3; d++) {
    G.lo[d] = c->sublo[d];
    G.nb[d] = static_cast<int>((c->subhi[d] - c->sublo[d]) * bininv);
    if (G.nb[d] == 0) G.nb[d] = 1;
    G.binv[d] = G.nb[d] / (c->subhi[d] - c->sublo[d]);
    nbins *= G.nb[d];
  }
  if (nbins == 1) return B200MD_OK;   // Atom::sort returns before permuting anything
  // scratch kept in the Langevin struct (grown on demand, freed by langevin_free): nothing is allocated, freed
  // or leaked per call, and no error path can drop it
  if (n > L.sort_cap) {
    if (L.sort_keys) CUDA_TRY(cudaFree(L.sort_keys));
    if (L.sort_vals) CUDA_TRY(cudaFree(L.sort_vals));
    L.sort_keys = nullptr; L.sort_vals = nullptr; L.sort_cap = 0;
    const size_t cap = (size_t)n + (size_t)n / 8 + 1024;
    CUDA_TRY(cudaMalloc(&L.sort_keys, sizeof(unsigned long long) * 2 * cap));
    CUDA_TRY(cudaMalloc(&L.sort_vals, sizeof(int) * 2 * cap));
    L.sort_cap = (int)cap;
  }
  unsigned long long *keys = L.sort_keys, *keys_out = L.sort_keys + L.sort_cap;
  int *vals = L.sort_vals, *vals_out = L.sort_vals + L.sort_cap;
  // only the bits the keys use: key < nbins * n
  int end_bit = 1;
  while (end_bit < 64 && ((unsigned long long)nbins * (unsigned long long)n) >> end_bit) end_bit++;
  size_t tmp_bytes = 0;
  CUDA_TRY(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys_out, vals, vals_out, n, 0, end_bit, c->stream));
  if (tmp_bytes > L.sort_tmp_bytes) {
    if (L.sort_tmp) CUDA_TRY(cudaFree(L.sort_tmp));
    L.sort_tmp = nullptr; L.sort_tmp_bytes = 0;
    CUDA_TRY(cudaMalloc(&L.sort_tmp, tmp_bytes + tmp_bytes / 8));
    L.sort_tmp_bytes = tmp_bytes + tmp_bytes / 8;
  }
  k_resort_keys<<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->xq, c->tag, L.rank_by_tag, G, (unsigned long long)n, keys, vals);
  CUDA_TRY(cudaGetLastError());
  tmp_bytes = L.sort_tmp_bytes;
  CUDA_TRY(cub::DeviceRadixSort::SortPairs(L.sort_tmp, tmp_bytes, keys, keys_out, vals, vals_out, n, 0, end_bit, c->stream));
  k_resort_assign<<<(n + 255) / 256, 256, 0, c->stream>>>(n, vals_out, L.rank_by_tag);
  CUDA_TRY(cudaGetLastError());
  c->launches += 3;
  return B200MD_OK;   // stream-ordered: the thermostat kernel that reads rank_by_tag follows on the same stream
}

int langevin_config(b200md_ctx *c, double t_start, double t_stop, double t_period, int seed, const double *ratio, double dt) {
  Langevin &L = c->lang;
  if (c->nranks != 1) {
    b200md_set_error("b200md: fix langevin runs on one rank (its random stream follows the host's local atom order)");
    return B200MD_ERR;
  }
  if (!c->host_order) { b200md_set_error("b200md: fix langevin must be set after the atoms and before the set-up"); return B200MD_ERR; }
  if (t_period <= 0.0) { b200md_set_error("Fix langevin period must be > 0.0"); return B200MD_ERR; }
  if (dt <= 0.0) { b200md_set_error("b200md: fix langevin needs the timestep"); return B200MD_ERR; }
  TRY(ranmars_device_init(c, seed));
  TRY(langevin_rank_from_host_order(c));
  L.t_start = t_start; L.t_stop = t_stop; L.t_period = t_period; L.dt = 0.0;
  L.ratio.assign(c->ntypes + 1, 1.0);
  if (ratio) for (int t = 1; t <= c->ntypes; t++) {
    if (ratio[t] <= 0.0) { b200md_set_error("Illegal fix langevin command"); return B200MD_ERR; }
    L.ratio[t] = ratio[t];
  }
  L.window = false;
  L.begin = L.end = L.now = c->step;
  L.on = true;
  return langevin_set_dt(c, dt);
}

// FixLangevin::init / reset_dt: the per-type prefactors (same operation order)
int langevin_set_dt(b200md_ctx *c, double dt) {
  Langevin &L = c->lang;
  if (dt == L.dt && L.gfac) return B200MD_OK;
  std::vector<double> g(2 * ((size_t)c->ntypes + 1), 0.0);
  for (int t = 1; t <= c->ntypes; t++) {
    double g1 = -c->h_mass[t] / L.t_period / c->ftm2v;
    double g2 = sqrt(c->h_mass[t]) / c->ftm2v;
    g2 *= sqrt(24.0 * c->boltz / L.t_period / dt / c->mvv2e);
    g1 *= 1.0 / L.ratio[t];
    g2 *= 1.0 / sqrt(L.ratio[t]);
    g[2 * t] = g1; g[2 * t + 1] = g2;
  }
  if (!L.gfac || L.gfac_n != (int)g.size()) {
    if (L.gfac) CUDA_TRY(cudaFree(L.gfac));
    CUDA_TRY(cudaMalloc(&L.gfac, sizeof(double) * g.size()));
    L.gfac_n = (int)g.size();
  }
  // stream-ordered against kernels still reading the old values
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  CUDA_TRY(cudaMemcpy(L.gfac, g.data(), sizeof(double) * g.size(), cudaMemcpyHostToDevice));
  L.dt = dt;
  return B200MD_OK;
}

// Modify::post_force -> FixLangevin::post_force at timestep L.now; with_final also does
// FixNVE::final_integrate in the same pass; with_next (needs with_final) the next step's initial_integrate, too
int langevin_post_force(b200md_ctx *c, bool with_final, double dtf, bool with_next, double dtv, bool check_dist) {
  Langevin &L = c->lang;
  const int n = c->nlocal;
  if (n != L.nrank) {
    b200md_set_error("b200md: fix langevin: the number of atoms changed since it was set");
    return B200MD_ERR;
  }
  double delta = (double)(L.now - L.begin);
  if (delta != 0.0) delta /= (double)(L.end - L.begin);
  const double t_target = L.t_start + delta * (L.t_stop - L.t_start);
  if (t_target < 0.0) { b200md_set_error("Fix langevin variable returned negative temperature"); return B200MD_ERR; }
  const double tsqrt = sqrt(t_target);
  TRY(ranmars_fill(c, 3 * n));
  if (n == 0) return B200MD_OK;
  const int nb = (n + 255) / 256;
#define B200_LANGEVIN(F, N, C)                                                                                              \
  k_langevin<F, N, C><<<nb, 256, 0, c->stream>>>(n, c->nmax, c->xq, c->v, c->f, c->tag, L.rank_by_tag, L.rnd, L.gfac, c->mass, \
                                                 tsqrt, dtf, dtv, c->xhold, c->triggersq, c->d_flags)
  if (with_final && with_next) { if (check_dist) B200_LANGEVIN(true, true, true); else B200_LANGEVIN(true, true, false); }
  else if (with_final) B200_LANGEVIN(true, false, false);
  else B200_LANGEVIN(false, false, false);
#undef B200_LANGEVIN
  CUDA_TRY(cudaGetLastError());
  c->launches++;
  return B200MD_OK;
}

// b200md_ranmars_device_stream: `fills` consecutive fills of n draws, as doubles
int langevin_test_stream(b200md_ctx *c, int seed, int n, int fills, double *out) {
  Langevin saved = c->lang;
  c->lang = Langevin();
  int rc = ranmars_device_init(c, seed);
  std::vector<uint32_t> h(n > 0 ? n : 1);
  for (int k = 0; rc == B200MD_OK && k < fills; k++) {
    rc = ranmars_fill(c, n);
    if (rc != B200MD_OK) break;
    if (cudaStreamSynchronize(c->stream) != cudaSuccess ||
        (n && cudaMemcpy(h.data(), c->lang.rnd, sizeof(uint32_t) * n, cudaMemcpyDeviceToHost) != cudaSuccess)) {
      b200md_set_error("b200md: CUDA error in the RanMars stream kernel");
      rc = B200MD_ERR;
      break;
    }
    for (int i = 0; i < n; i++) out[(size_t)k * n + i] = (double)h[i] * (1.0 / 16777216.0);
  }
  langevin_free(c);
  c->lang = saved;
  return rc;
}
