// comm.cu — ghost-atom halo exchange and atom migration on packed device buffers.
//
// Reference behaviour followed (/root/reference/src):
//   CommBrick::setup          comm_brick.cpp:191-483   -> comm_setup (schedule from host.cpp)
//   CommBrick::borders        comm_brick.cpp:771-979   -> comm_borders
//   AtomVec::pack_border      atom_vec.cpp:794-830     -> k_pack_border / k_self_border (x + pbc*prd)
//   CommBrick::forward_comm   comm_brick.cpp:538-591   -> comm_forward
//   CommBrick::exchange       comm_brick.cpp:652-759   -> comm_exchange
//   CommBrick::forward_comm(Pair*) :986-1016           -> comm_forward_scalar (EAM fp)
//   CommBrick::reverse_comm   :598-639, (Pair*) :1023-1053 -> comm_reverse_force / _scalar
//   MPI_Allreduce sites (neighbor.cpp:2387, compute_*.cpp)  -> comm_allreduce_*
// Transport: a swap whose partner is this rank is a device gather kernel; remote swaps are
// grouped ncclSend/ncclRecv of packed double4 records over NVLink, received in place into
// the ghost slots (ghosts of one swap are contiguous), both directions of a dimension in
// one NCCL group.  One process per GPU; the unique id is distributed by the launcher.
// B200MD_HALO=direct (opt-in): the staged borders additionally carry each ghost's owner (rank, index,
// shift); per-step forward / reverse traffic then goes owner <-> holder directly in ONE NCCL group
// (a message per peer rank) instead of three dependent stages — see "direct halo" below.
#include "comm.h"

extern "C" int b200md_comm_unique_id(void *id128) {
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId size");
  ncclUniqueId id;
  NCCL_TRY(ncclGetUniqueId(&id));
  memcpy(id128, &id, 128);
  return B200MD_OK;
}

extern "C" int b200md_comm_init(b200md_ctx *c, const void *id128) {
  if (c->nranks == 1) return B200MD_OK;
  CUDA_TRY(cudaSetDevice(c->device));
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  c->nccl = new NcclState();
  NCCL_TRY(ncclCommInitRank(&c->nccl->comm, c->nranks, id, c->rank));
  CUDA_TRY(cudaMalloc(&c->nccl->d_counts, 8 * sizeof(int)));
  CUDA_TRY(cudaMallocHost(&c->nccl->h_counts, 8 * sizeof(int)));
  CUDA_TRY(cudaMalloc(&c->nccl->d_red, 32 * sizeof(double)));
  return B200MD_OK;
}

void comm_free(b200md_ctx *c) {
  h26_free(c);
  for (auto &s : c->swaps) if (s.sendlist) cudaFree(s.sendlist);
  c->swaps.clear();
  DirectHalo &D = c->direct;
  if (D.req_idx) cudaFree(D.req_idx);
  if (D.ghost_slot) cudaFree(D.ghost_slot);
  if (D.serve_idx) cudaFree(D.serve_idx);
  if (D.d_cnt) cudaFree(D.d_cnt);
  if (D.h_cnt) cudaFreeHost(D.h_cnt);
  D.req_idx = D.ghost_slot = D.serve_idx = D.d_cnt = D.h_cnt = nullptr;
  D.req_cap = D.serve_cap = 0;
  D.ready = false;
  if (c->nccl) {
    if (c->nccl->comm) ncclCommDestroy(c->nccl->comm);
    if (c->nccl->d_counts) cudaFree(c->nccl->d_counts);
    if (c->nccl->h_counts) cudaFreeHost(c->nccl->h_counts);
    if (c->nccl->d_red) cudaFree(c->nccl->d_red);
    delete c->nccl;
    c->nccl = nullptr;
  }
}

int comm_allreduce_sum(b200md_ctx *c, double *v, int n) {
  if (c->nranks == 1) return B200MD_OK;
  if (!c->nccl) { b200md_set_error("b200md: comm not initialised (b200md_comm_init)"); return B200MD_ERR; }
  CUDA_TRY(cudaMemcpyAsync(c->nccl->d_red, v, n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
  NCCL_TRY(ncclAllReduce(c->nccl->d_red, c->nccl->d_red, n, ncclDouble, ncclSum, c->nccl->comm, c->stream));
  CUDA_TRY(cudaMemcpyAsync(v, c->nccl->d_red, n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

int comm_allreduce_max_int(b200md_ctx *c, int *v) {
  if (c->nranks == 1) return B200MD_OK;
  if (!c->nccl) { b200md_set_error("b200md: comm not initialised (b200md_comm_init)"); return B200MD_ERR; }
  CUDA_TRY(cudaMemcpyAsync(c->nccl->d_counts, v, sizeof(int), cudaMemcpyHostToDevice, c->stream));
  NCCL_TRY(ncclAllReduce(c->nccl->d_counts, c->nccl->d_counts, 1, ncclInt, ncclMax, c->nccl->comm, c->stream));
  CUDA_TRY(cudaMemcpyAsync(v, c->nccl->d_counts, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

int comm_setup(b200md_ctx *c) {
  TRY(b200md_comm_schedule(c->boxlo, c->boxhi, c->periodic, c->procgrid, c->myloc, c->cutghost, &c->sched));
  for (auto &s : c->swaps) if (s.sendlist) cudaFree(s.sendlist);
  c->swaps.assign(c->sched.nswap, Swap());
  for (int i = 0; i < c->sched.nswap; i++) {
    Swap &s = c->swaps[i];
    s.dim = c->sched.dim[i]; s.ineed = c->sched.ineed[i]; s.sendflag = c->sched.sendflag[i];
    s.sendproc = c->sched.sendproc[i]; s.recvproc = c->sched.recvproc[i];
    for (int d = 0; d < 3; d++) s.pbc[d] = c->sched.pbc[i][d];
    s.slablo = c->sched.slablo[i]; s.slabhi = c->sched.slabhi[i];
  }
  // halo mode on several ranks: direct (halo26.cu) when every dimension needs one ghost layer, else the owner-tracking
  // variant of the staged borders; "staged" keeps the reference's forwarding stages for the per-step traffic too
  TRY(h26_setup(c, c->halo_mode == 2));
  c->direct.on = c->nranks > 1 && (c->halo_mode == 1 || (c->halo_mode == 2 && !c->h26.on));
  c->direct.ready = false;
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------------------
__device__ __forceinline__ double comp(const double4 &q, int d) { return d == 0 ? q.x : (d == 1 ? q.y : q.z); }

__global__ void k_slab_flag(const double4 *__restrict__ xq, int nfirst, int n, int dim, double lo, double hi,
                            int *__restrict__ flag) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n) return;
  if (k == n) { flag[k] = 0; return; }
  const double x = comp(xq[nfirst + k], dim);
  flag[k] = (x >= lo && x <= hi) ? 1 : 0;  // comm_brick.cpp:838
}
__global__ void k_compact(const int *__restrict__ flag, const int *__restrict__ pos, int nfirst, int n,
                          int *__restrict__ list) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  if (flag[k]) list[pos[k]] = nfirst + k;
}

__device__ __forceinline__ int pbc_code_add(int code, int px, int py, int pz) {
  int cx = (code & 15) + px, cy = ((code >> 4) & 15) + py, cz = ((code >> 8) & 15) + pz;
  return cx | (cy << 4) | (cz << 8);
}
#define PBC_ZERO (8 | (8 << 4) | (8 << 8))


// self swap of borders: append ghosts, record flattened source
__global__ void k_self_border(int n, const int *__restrict__ list, int first, int nlocal, double4 *__restrict__ xq,
                              int *__restrict__ tag, int *__restrict__ ghost_src, int *__restrict__ ghost_pbc,
                              int px, int py, int pz, double prdx, double prdy, double prdz,
                              int *__restrict__ ghost_rank, int me) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int j = list[k];
  double4 q = xq[j];
  q.x = shift_n(q.x, px, prdx); q.y = shift_n(q.y, py, prdy); q.z = shift_n(q.z, pz, prdz);
  xq[first + k] = q;
  tag[first + k] = tag[j];
  int src = j, code = PBC_ZERO;
  if (j >= nlocal) { src = ghost_src[j]; code = ghost_pbc[j]; }
  ghost_src[first + k] = src;
  ghost_pbc[first + k] = pbc_code_add(code, px, py, pz);
  if (ghost_rank) ghost_rank[first + k] = j >= nlocal ? ghost_rank[j] : me;   // direct halo: the owner's rank
}
// direct halo: owner (index on its rank, accumulated shift, rank) of every atom of a remote border swap
__global__ void k_pack_border_owner(int n, const int *__restrict__ list, int nlocal, int me,
                                    const int *__restrict__ ghost_src, const int *__restrict__ ghost_pbc,
                                    const int *__restrict__ ghost_rank, int px, int py, int pz,
                                    int *__restrict__ o_src, int *__restrict__ o_code, int *__restrict__ o_rank) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int j = list[k];
  int src = j, code = PBC_ZERO, rk = me;
  if (j >= nlocal) { src = ghost_src[j]; code = ghost_pbc[j]; rk = ghost_rank[j]; }
  o_src[k] = src; o_code[k] = pbc_code_add(code, px, py, pz); o_rank[k] = rk;
}
// remote swap: packed records (positions with shift) + tags
__global__ void k_pack_border(int n, const int *__restrict__ list, const double4 *__restrict__ xq,
                              const int *__restrict__ tag, double4 *__restrict__ buf, int *__restrict__ tagbuf,
                              int px, int py, int pz, double prdx, double prdy, double prdz) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int j = list[k];
  double4 q = xq[j];
  q.x = shift_n(q.x, px, prdx); q.y = shift_n(q.y, py, prdy); q.z = shift_n(q.z, pz, prdz);
  buf[k] = q;
  if (tagbuf) tagbuf[k] = tag[j];
}
__global__ void k_self_forward(int n, const int *__restrict__ list, int first, double4 *__restrict__ xq,
                               int px, int py, int pz, double prdx, double prdy, double prdz) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  double4 q = xq[list[k]];
  q.x = shift_n(q.x, px, prdx); q.y = shift_n(q.y, py, prdy); q.z = shift_n(q.z, pz, prdz);
  xq[first + k] = q;
}
// all swaps are self: one gather from the owning atom
__global__ void k_forward_flat(int nghost, int nlocal, double4 *__restrict__ xq, const int *__restrict__ ghost_src,
                               const int *__restrict__ ghost_pbc, double prdx, double prdy, double prdz) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nghost) return;
  const int code = ghost_pbc[nlocal + g];
  double4 q = xq[ghost_src[nlocal + g]];
  q.x = shift_n(q.x, (code & 15) - 8, prdx);
  q.y = shift_n(q.y, ((code >> 4) & 15) - 8, prdy);
  q.z = shift_n(q.z, ((code >> 8) & 15) - 8, prdz);
  xq[nlocal + g] = q;
}
__global__ void k_scalar_flat(int nghost, int nlocal, double *__restrict__ a, const int *__restrict__ ghost_src) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nghost) return;
  a[nlocal + g] = a[ghost_src[nlocal + g]];
}
__global__ void k_scalar_self(int n, const int *__restrict__ list, int first, double *__restrict__ a) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  a[first + k] = a[list[k]];
}
__global__ void k_scalar_pack(int n, const int *__restrict__ list, const double *__restrict__ a, double *__restrict__ buf) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  buf[k] = a[list[k]];
}


// ---------------------------------------------------------------------------------------
// direct halo (B200MD_HALO=direct)
// The reference forwards ghosts dimension by dimension (x, then y including the x-ghosts, then z):
// 3 dependent stages of 2 messages, which on NVLink cost latency, not bandwidth.  Once the staged
// borders have run, each ghost's owner is known, so every later forward_comm / reverse_comm of the
// same list goes owner <-> holder in one hop: ghosts are grouped by owner rank (request lists are
// exchanged once per rebuild), and a step needs one pack kernel, ONE NCCL group with one send and
// one receive per peer rank, and one unpack kernel that applies the accumulated periodic shift
// exactly as the stages would have (one add of prd per crossed image and dimension).
// ---------------------------------------------------------------------------------------
__global__ void k_direct_count(int nghost, int nlocal, const int *__restrict__ ghost_rank, int *__restrict__ cnt) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nghost) return;
  atomicAdd(&cnt[ghost_rank[nlocal + g]], 1);
}
__global__ void k_direct_scatter(int nghost, int nlocal, const int *__restrict__ ghost_rank, const int *__restrict__ ghost_src,
                                 int *__restrict__ cursor, int *__restrict__ req_idx, int *__restrict__ ghost_slot) {
  int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nghost) return;
  const int pos = atomicAdd(&cursor[ghost_rank[nlocal + g]], 1);
  req_idx[pos] = ghost_src[nlocal + g];
  ghost_slot[pos] = nlocal + g;
}
template <bool SCALAR>
__global__ void k_direct_pack(int n, const int *__restrict__ serve_idx, const double4 *__restrict__ xq,
                              const double *__restrict__ a, double *__restrict__ buf) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  if (SCALAR) buf[k] = a[serve_idx[k]];
  else reinterpret_cast<double4 *>(buf)[k] = xq[serve_idx[k]];
}
// requests [0, nremote) were answered through buf, [nremote, n) are images of my own atoms
template <bool SCALAR>
__global__ void k_direct_unpack(int n, int nremote, const int *__restrict__ ghost_slot, const int *__restrict__ req_idx,
                                const double *__restrict__ buf, double4 *__restrict__ xq, double *__restrict__ a,
                                const int *__restrict__ ghost_pbc, double prdx, double prdy, double prdz) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int g = ghost_slot[k];
  if (SCALAR) { a[g] = k < nremote ? buf[k] : a[req_idx[k]]; return; }
  double4 q = k < nremote ? reinterpret_cast<const double4 *>(buf)[k] : xq[req_idx[k]];
  const int code = ghost_pbc[g];
  q.x = shift_n(q.x, (code & 15) - 8, prdx);
  q.y = shift_n(q.y, ((code >> 4) & 15) - 8, prdy);
  q.z = shift_n(q.z, ((code >> 8) & 15) - 8, prdz);
  xq[g] = q;
}
template <int NV>
__global__ void k_direct_rev_pack(int n, int nremote, int nmax, const int *__restrict__ ghost_slot,
                                  const int *__restrict__ req_idx, double *__restrict__ a, double *__restrict__ buf) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int g = ghost_slot[k];
#pragma unroll
  for (int d = 0; d < NV; d++) {
    const double v = a[g + d * (size_t)nmax];
    if (k < nremote) buf[(size_t)k * NV + d] = v;
    else atomicAdd(&a[req_idx[k] + d * (size_t)nmax], v);
  }
}
template <int NV>
__global__ void k_direct_rev_unpack(int n, int nmax, const int *__restrict__ serve_idx, double *__restrict__ a,
                                    const double *__restrict__ buf) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int j = serve_idx[k];
#pragma unroll
  for (int d = 0; d < NV; d++) atomicAdd(&a[j + d * (size_t)nmax], buf[(size_t)k * NV + d]);
}

static int direct_setup(b200md_ctx *c) {
  DirectHalo &D = c->direct;
  NcclState *nc = c->nccl;
  if (!nc) { b200md_set_error("b200md: comm not initialised (b200md_comm_init)"); return B200MD_ERR; }
  cudaStream_t st = c->stream;
  const int nr = c->nranks, me = c->rank, ng = c->nghost;
  if (!D.d_cnt) {
    CUDA_TRY(cudaMalloc(&D.d_cnt, sizeof(int) * (2 * nr + (size_t)nr * nr)));
    CUDA_TRY(cudaMallocHost(&D.h_cnt, sizeof(int) * (2 * nr + (size_t)nr * nr)));
  }
  int *d_row = D.d_cnt, *d_cursor = D.d_cnt + nr, *d_mat = D.d_cnt + 2 * nr;
  int *h_row = D.h_cnt, *h_cursor = D.h_cnt + nr, *h_mat = D.h_cnt + 2 * nr;
  // my requests per owner rank, and everybody's (row p of the matrix = rank p's requests)
  CUDA_TRY(cudaMemsetAsync(d_row, 0, sizeof(int) * nr, st));
  if (ng) k_direct_count<<<nblk(ng), 256, 0, st>>>(ng, c->nlocal, c->ghost_rank, d_row);
  NCCL_TRY(ncclAllGather(d_row, d_mat, nr, ncclInt, nc->comm, st));
  CUDA_TRY(cudaMemcpyAsync(h_row, d_row, sizeof(int) * nr, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaMemcpyAsync(h_mat, d_mat, sizeof(int) * (size_t)nr * nr, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  D.req_cnt.assign(h_row, h_row + nr);
  D.req_off.assign(nr, 0); D.serve_cnt.assign(nr, 0); D.serve_off.assign(nr, 0);
  int off = 0, soff = 0;
  for (int p = 0; p < nr; p++) {
    if (p == me) continue;
    D.req_off[p] = off; off += D.req_cnt[p];
    D.serve_cnt[p] = h_mat[(size_t)p * nr + me];
    D.serve_off[p] = soff; soff += D.serve_cnt[p];
  }
  D.nreq_remote = off;
  D.req_off[me] = off; off += D.req_cnt[me];   // images of my own atoms last
  D.nreq = off;
  D.nserve = soff;
  if (D.nreq != ng) { b200md_set_error("b200md: direct halo: a ghost without a valid owner rank"); return B200MD_ERR; }
  if (D.nreq > D.req_cap) {
    if (D.req_idx) CUDA_TRY(cudaFree(D.req_idx));
    if (D.ghost_slot) CUDA_TRY(cudaFree(D.ghost_slot));
    D.req_cap = D.nreq + D.nreq / 4 + 1024;
    CUDA_TRY(cudaMalloc(&D.req_idx, sizeof(int) * D.req_cap));
    CUDA_TRY(cudaMalloc(&D.ghost_slot, sizeof(int) * D.req_cap));
  }
  if (D.nserve > D.serve_cap) {
    if (D.serve_idx) CUDA_TRY(cudaFree(D.serve_idx));
    D.serve_cap = D.nserve + D.nserve / 4 + 1024;
    CUDA_TRY(cudaMalloc(&D.serve_idx, sizeof(int) * D.serve_cap));
  }
  for (int p = 0; p < nr; p++) h_cursor[p] = D.req_off[p];
  CUDA_TRY(cudaMemcpyAsync(d_cursor, h_cursor, sizeof(int) * nr, cudaMemcpyHostToDevice, st));
  if (ng) k_direct_scatter<<<nblk(ng), 256, 0, st>>>(ng, c->nlocal, c->ghost_rank, c->ghost_src, d_cursor, D.req_idx, D.ghost_slot);
  c->launches += 2;
  // the request lists travel to the owners once per rebuild
  NCCL_TRY(ncclGroupStart());
  for (int p = 0; p < nr; p++) {
    if (p == me) continue;
    if (D.req_cnt[p]) NCCL_TRY(ncclSend(D.req_idx + D.req_off[p], D.req_cnt[p], ncclInt, p, nc->comm, st));
    if (D.serve_cnt[p]) NCCL_TRY(ncclRecv(D.serve_idx + D.serve_off[p], D.serve_cnt[p], ncclInt, p, nc->comm, st));
  }
  NCCL_TRY(ncclGroupEnd());
  CUDA_TRY(cudaGetLastError());
  D.ready = true;
  return B200MD_OK;
}

// owners -> holders: positions (shift applied by the holder) or one per-atom scalar
template <bool SCALAR>
static int direct_forward(b200md_ctx *c, double *arr) {
  DirectHalo &D = c->direct;
  NcclState *nc = c->nccl;
  cudaStream_t st = c->stream;
  const int rec = SCALAR ? 1 : 4, nr = c->nranks, me = c->rank;
  TRY(ensure_buf(c->sendbuf, c->sendcap, (size_t)D.nserve * rec + 8));
  TRY(ensure_buf(c->recvbuf, c->recvcap, (size_t)D.nreq_remote * rec + 8));
  if (D.nserve) { k_direct_pack<SCALAR><<<nblk(D.nserve), 256, 0, st>>>(D.nserve, D.serve_idx, c->xq, arr, c->sendbuf); c->launches++; }
  NCCL_TRY(ncclGroupStart());
  for (int p = 0; p < nr; p++) {
    if (p == me) continue;
    if (D.serve_cnt[p]) NCCL_TRY(ncclSend(c->sendbuf + (size_t)D.serve_off[p] * rec, (size_t)D.serve_cnt[p] * rec, ncclDouble, p, nc->comm, st));
    if (D.req_cnt[p]) NCCL_TRY(ncclRecv(c->recvbuf + (size_t)D.req_off[p] * rec, (size_t)D.req_cnt[p] * rec, ncclDouble, p, nc->comm, st));
  }
  NCCL_TRY(ncclGroupEnd());
  if (D.nreq) {
    k_direct_unpack<SCALAR><<<nblk(D.nreq), 256, 0, st>>>(D.nreq, D.nreq_remote, D.ghost_slot, D.req_idx, c->recvbuf, c->xq, arr,
                                                         c->ghost_pbc, c->prd[0], c->prd[1], c->prd[2]);
    c->launches++;
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

// holders -> owners: every ghost copy adds its NV values straight into its owner (FP64 RED)
template <int NV>
static int direct_reverse(b200md_ctx *c, double *arr) {
  DirectHalo &D = c->direct;
  NcclState *nc = c->nccl;
  cudaStream_t st = c->stream;
  const int nr = c->nranks, me = c->rank;
  TRY(ensure_buf(c->sendbuf, c->sendcap, (size_t)D.nreq_remote * NV + 8));
  TRY(ensure_buf(c->recvbuf, c->recvcap, (size_t)D.nserve * NV + 8));
  if (D.nreq) { k_direct_rev_pack<NV><<<nblk(D.nreq), 256, 0, st>>>(D.nreq, D.nreq_remote, c->nmax, D.ghost_slot, D.req_idx, arr, c->sendbuf); c->launches++; }
  NCCL_TRY(ncclGroupStart());
  for (int p = 0; p < nr; p++) {
    if (p == me) continue;
    if (D.req_cnt[p]) NCCL_TRY(ncclSend(c->sendbuf + (size_t)D.req_off[p] * NV, (size_t)D.req_cnt[p] * NV, ncclDouble, p, nc->comm, st));
    if (D.serve_cnt[p]) NCCL_TRY(ncclRecv(c->recvbuf + (size_t)D.serve_off[p] * NV, (size_t)D.serve_cnt[p] * NV, ncclDouble, p, nc->comm, st));
  }
  NCCL_TRY(ncclGroupEnd());
  if (D.nserve) { k_direct_rev_unpack<NV><<<nblk(D.nserve), 256, 0, st>>>(D.nserve, c->nmax, D.serve_idx, arr, c->recvbuf); c->launches++; }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// borders
// ---------------------------------------------------------------------------------------
// slab selection of the two swaps of a pair (the - and + direction of one dimension scan the same
// atom range): count per 1024-atom tile, offsets by one small scan, ordered compaction.  One pass
// over the positions for the counts, one for the compaction; the send lists come out ascending
// like the reference's loop (comm_brick.cpp:836-842), so the ghost order is deterministic.
#define SLAB_T 256
#define SLAB_ITEMS 4
#define SLAB_TILE (SLAB_T * SLAB_ITEMS)
struct SlabPair { double lo[2], hi[2]; int use[2]; };

__device__ __forceinline__ int slab_bits(const double4 *__restrict__ xq, int idx, int dim, const SlabPair &P) {
  const double x = comp(xq[idx], dim);
  int b = 0;
  if (P.use[0] && x >= P.lo[0] && x <= P.hi[0]) b |= 1;   // comm_brick.cpp:838
  if (P.use[1] && x >= P.lo[1] && x <= P.hi[1]) b |= 2;
  return b;
}
__global__ void __launch_bounds__(SLAB_T) k_slab_count2(const double4 *__restrict__ xq, int nfirst, int n, int dim, SlabPair P,
                                                        int *__restrict__ tile_cnt, int ntiles) {
  __shared__ int s_cnt[2];
  if (threadIdx.x < 2) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  int ca = 0, cb = 0;
  const int base = blockIdx.x * SLAB_TILE;
#pragma unroll
  for (int it = 0; it < SLAB_ITEMS; it++) {
    const int k = base + it * SLAB_T + threadIdx.x;
    if (k < n) { const int b = slab_bits(xq, nfirst + k, dim, P); ca += b & 1; cb += b >> 1; }
  }
  ca = __reduce_add_sync(0xffffffffu, ca); cb = __reduce_add_sync(0xffffffffu, cb);
  if ((threadIdx.x & 31) == 0) { if (ca) atomicAdd(&s_cnt[0], ca); if (cb) atomicAdd(&s_cnt[1], cb); }
  __syncthreads();
  if (threadIdx.x < 2) tile_cnt[threadIdx.x * ntiles + blockIdx.x] = s_cnt[threadIdx.x];
}
// one block: exclusive scan of both count rows in place, totals to totals[0..1]
__global__ void __launch_bounds__(SLAB_T) k_slab_offsets(int *tile_cnt, int ntiles, int *totals) {
  __shared__ int s_warp[33];
  for (int row = 0; row < 2; row++) {
    int carry = 0;
    int *t = tile_cnt + row * ntiles;
    for (int base = 0; base < ntiles; base += SLAB_T) {
      const int i = base + threadIdx.x;
      const int v = i < ntiles ? t[i] : 0;
      // block exclusive scan
      const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
      int inc = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
      if (lane == 31) s_warp[wid] = inc;
      __syncthreads();
      if (wid == 0) {
        int w = lane < SLAB_T / 32 ? s_warp[lane] : 0, winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += y; }
        if (lane < SLAB_T / 32) s_warp[lane] = winc - w;
        if (lane == SLAB_T / 32 - 1) s_warp[32] = winc;
      }
      __syncthreads();
      if (i < ntiles) t[i] = carry + s_warp[wid] + inc - v;
      carry += s_warp[32];
      __syncthreads();
    }
    if (threadIdx.x == 0) totals[row] = carry;
  }
}
__global__ void __launch_bounds__(SLAB_T) k_slab_compact2(const double4 *__restrict__ xq, int nfirst, int n, int dim, SlabPair P,
                                                          const int *__restrict__ tile_off, int ntiles,
                                                          int *__restrict__ list_a, int *__restrict__ list_b) {
  __shared__ int s_w[2][SLAB_T / 32];
  __shared__ int s_run[2];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (threadIdx.x < 2) s_run[threadIdx.x] = tile_off[threadIdx.x * ntiles + blockIdx.x];
  __syncthreads();
  const int base = blockIdx.x * SLAB_TILE;
  for (int it = 0; it < SLAB_ITEMS; it++) {
    const int k = base + it * SLAB_T + threadIdx.x;
    const int b = k < n ? slab_bits(xq, nfirst + k, dim, P) : 0;
    const unsigned ma = __ballot_sync(0xffffffffu, b & 1), mb = __ballot_sync(0xffffffffu, b & 2);
    if (lane == 0) { s_w[0][wid] = __popc(ma); s_w[1][wid] = __popc(mb); }
    __syncthreads();
    int wa = 0, wb = 0, ta = 0, tb = 0;
#pragma unroll
    for (int w = 0; w < SLAB_T / 32; w++) { if (w < wid) { wa += s_w[0][w]; wb += s_w[1][w]; } ta += s_w[0][w]; tb += s_w[1][w]; }
    const unsigned below = (1u << lane) - 1u;
    if (b & 1) list_a[s_run[0] + wa + __popc(ma & below)] = nfirst + k;
    if (b & 2) list_b[s_run[1] + wb + __popc(mb & below)] = nfirst + k;
    __syncthreads();
    if (threadIdx.x == 0) { s_run[0] += ta; s_run[1] += tb; }
    __syncthreads();
  }
}

static int grow_sendlist(Swap &s, int nsend) {
  if (nsend > s.sendcap) {
    if (s.sendlist) CUDA_TRY(cudaFree(s.sendlist));
    s.sendcap = nsend + nsend / 4 + 256;
    CUDA_TRY(cudaMalloc(&s.sendlist, sizeof(int) * s.sendcap));
  }
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// One rank, one ghost layer per side: all borders in one pass ("self26").  Every swap partner is this rank, so the
// staged ghost set is the product set over the dimensions of the per-dimension send slabs (halo26.cu's argument): an
// owned atom has an image in direction (dx,dy,dz) != 0 iff for every non-zero component it lies in that dimension's
// slab.  Three kernels and ONE host sync per rebuild instead of the staged 9 selection kernels + 6 copies + 3 syncs:
// count per (direction, tile), scan per direction, fill.  Ghost order: direction-major, ascending owner index inside a
// direction (deterministic).  Image position = x + pbc*prd, one addition per crossed dimension (AtomVec::pack_border).
// ---------------------------------------------------------------------------------------
#define S26_T 256
#define S26_ITEMS 4
#define S26_TILE (S26_T * S26_ITEMS)
struct S26Geom {
  double lhi[3], rlo[3];     // left swap: x <= lhi; right swap: x >= rlo (first swap of a dimension: one-sided)
  double llo[3], rhi[3];
  int useL[3], useR[3];
  int pbcL[3], pbcR[3];      // periodic shift multiple applied by that swap in its own dimension
};
__device__ __forceinline__ unsigned s26_dirs(const double4 q, const S26Geom &G) {
  bool L[3], R[3];
  const double x[3] = {q.x, q.y, q.z};
#pragma unroll
  for (int k = 0; k < 3; k++) {
    L[k] = G.useL[k] && x[k] >= G.llo[k] && x[k] <= G.lhi[k];
    R[k] = G.useR[k] && x[k] >= G.rlo[k] && x[k] <= G.rhi[k];
  }
  unsigned m = 0;
#pragma unroll
  for (int d = 0; d < 27; d++) {
    if (d == 13) continue;
    const int c0 = d % 3 - 1, c1 = (d / 3) % 3 - 1, c2 = d / 9 - 1;
    const bool ok = (c0 == 0 || (c0 < 0 ? L[0] : R[0])) && (c1 == 0 || (c1 < 0 ? L[1] : R[1])) && (c2 == 0 || (c2 < 0 ? L[2] : R[2]));
    m |= (ok ? 1u : 0u) << d;
  }
  return m;
}
__global__ void __launch_bounds__(S26_T) k_s26_count(const double4 *__restrict__ xq, int n, S26Geom G, int *__restrict__ tile_cnt,
                                                     int ntiles) {
  __shared__ int s_cnt[27];
  if (threadIdx.x < 27) s_cnt[threadIdx.x] = 0;
  __syncthreads();
  const int base = blockIdx.x * S26_TILE;
#pragma unroll
  for (int it = 0; it < S26_ITEMS; it++) {
    const int i = base + it * S26_T + threadIdx.x;
    unsigned m = i < n ? s26_dirs(xq[i], G) : 0u;
    if (__any_sync(0xffffffffu, m != 0)) {
      for (int d = 0; d < 27; d++) {
        const unsigned b = __ballot_sync(0xffffffffu, (m >> d) & 1u);
        if (b && (threadIdx.x & 31) == 0) atomicAdd(&s_cnt[d], __popc(b));
      }
    }
  }
  __syncthreads();
  if (threadIdx.x < 27) tile_cnt[threadIdx.x * ntiles + blockIdx.x] = s_cnt[threadIdx.x];
}
// one block per direction: exclusive scan of its tile counts in place, total to totals[d]
__global__ void __launch_bounds__(S26_T) k_s26_offsets(int *tile_cnt, int ntiles, int *totals) {
  __shared__ int s_warp[33];
  int *t = tile_cnt + blockIdx.x * ntiles;
  int carry = 0;
  for (int base = 0; base < ntiles; base += S26_T) {
    const int i = base + threadIdx.x;
    const int v = i < ntiles ? t[i] : 0;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += y; }
    if (lane == 31) s_warp[wid] = inc;
    __syncthreads();
    if (wid == 0) {
      int w = lane < S26_T / 32 ? s_warp[lane] : 0, winc = w;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, winc, o); if (lane >= o) winc += y; }
      s_warp[lane] = winc - w;
      if (lane == 31) s_warp[32] = winc;
    }
    __syncthreads();
    if (i < ntiles) t[i] = carry + s_warp[wid] + inc - v;
    carry += s_warp[32];
    __syncthreads();
  }
  if (threadIdx.x == 0) totals[blockIdx.x] = carry;
}
struct S26Base { int v[27]; };
__global__ void __launch_bounds__(S26_T) k_s26_fill(double4 *__restrict__ xq, int *__restrict__ tag, int n, S26Geom G,
                                                    const int *__restrict__ tile_off, int ntiles, S26Base first,
                                                    int *__restrict__ ghost_src, int *__restrict__ ghost_pbc, double prdx,
                                                    double prdy, double prdz) {
  // per direction, per (item, warp): how many images come before -> ascending owner index inside the tile
  __shared__ int s_cnt[27][S26_ITEMS * (S26_T / 32)];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int base = blockIdx.x * S26_TILE;
  unsigned m[S26_ITEMS];
  for (int k = threadIdx.x; k < 27 * S26_ITEMS * (S26_T / 32); k += S26_T) (&s_cnt[0][0])[k] = 0;
  __syncthreads();
  bool any_tile = false;
#pragma unroll
  for (int it = 0; it < S26_ITEMS; it++) {
    const int i = base + it * S26_T + threadIdx.x;
    m[it] = i < n ? s26_dirs(xq[i], G) : 0u;
    if (__any_sync(0xffffffffu, m[it] != 0)) {
      any_tile = true;
      for (int d = 0; d < 27; d++) {
        const unsigned b = __ballot_sync(0xffffffffu, (m[it] >> d) & 1u);
        if (b && lane == 0) s_cnt[d][it * (S26_T / 32) + wid] = __popc(b);
      }
    }
  }
  if (!__syncthreads_or(any_tile)) return;
  if (threadIdx.x < 27) {   // exclusive scan over the (item, warp) slots of one direction
    int run = 0;
    for (int k = 0; k < S26_ITEMS * (S26_T / 32); k++) { const int v = s_cnt[threadIdx.x][k]; s_cnt[threadIdx.x][k] = run; run += v; }
  }
  __syncthreads();
#pragma unroll
  for (int it = 0; it < S26_ITEMS; it++) {
    if (!__any_sync(0xffffffffu, m[it] != 0)) continue;
    const int i = base + it * S26_T + threadIdx.x;
    const double4 q = i < n ? xq[i] : make_double4(0, 0, 0, 0);
    for (int d = 0; d < 27; d++) {
      const unsigned b = __ballot_sync(0xffffffffu, (m[it] >> d) & 1u);
      if (!((m[it] >> d) & 1u)) continue;
      const int slot = first.v[d] + tile_off[d * ntiles + blockIdx.x] + s_cnt[d][it * (S26_T / 32) + wid] + __popc(b & ((1u << lane) - 1u));
      const int c0 = d % 3 - 1, c1 = (d / 3) % 3 - 1, c2 = d / 9 - 1;
      const int px = c0 < 0 ? G.pbcL[0] : (c0 > 0 ? G.pbcR[0] : 0);
      const int py = c1 < 0 ? G.pbcL[1] : (c1 > 0 ? G.pbcR[1] : 0);
      const int pz = c2 < 0 ? G.pbcL[2] : (c2 > 0 ? G.pbcR[2] : 0);
      double4 g = q;
      g.x = shift_n(q.x, px, prdx); g.y = shift_n(q.y, py, prdy); g.z = shift_n(q.z, pz, prdz);
      xq[slot] = g;
      tag[slot] = tag[i];
      ghost_src[slot] = i;
      ghost_pbc[slot] = pbc_code_add(PBC_ZERO, px, py, pz);
    }
  }
}

static bool self26_usable(const b200md_ctx *c) {
  if (c->nranks != 1 || !c->self26) return false;
  for (int d = 0; d < 3; d++) if (c->sched.maxneed[d] > 1) return false;
  for (auto &s : c->swaps) if (s.sendproc != c->rank || s.recvproc != c->rank) return false;
  return true;
}

static int comm_borders_self26(b200md_ctx *c) {
  cudaStream_t st = c->stream;
  const int n = c->nlocal;
  c->nghost = 0;
  c->all_self = true;
  c->direct.ready = false;
  for (auto &s : c->swaps) { s.sendnum = 0; s.recvnum = 0; s.firstrecv = n; }
  S26Geom G;
  int iswap = 0;
  for (int dim = 0; dim < 3; dim++) {
    G.useL[dim] = G.useR[dim] = 0; G.pbcL[dim] = G.pbcR[dim] = 0;
    G.llo[dim] = G.rlo[dim] = 0; G.lhi[dim] = G.rhi[dim] = 0;
    if (c->sched.maxneed[dim] == 1) {
      const Swap &a = c->swaps[iswap], &b = c->swaps[iswap + 1];
      G.useL[dim] = a.sendflag ? 1 : 0; G.llo[dim] = a.slablo; G.lhi[dim] = a.slabhi; G.pbcL[dim] = a.pbc[dim];
      G.useR[dim] = b.sendflag ? 1 : 0; G.rlo[dim] = b.slablo; G.rhi[dim] = b.slabhi; G.pbcR[dim] = b.pbc[dim];
      iswap += 2;
    }
  }
  if (n == 0) return B200MD_OK;
  const int ntiles = (n + S26_TILE - 1) / S26_TILE;
  if (27 * ntiles + 64 > c->scan_cap) {
    if (c->scan_block) CUDA_TRY(cudaFree(c->scan_block));
    c->scan_cap = 27 * ntiles + 1024;
    CUDA_TRY(cudaMalloc(&c->scan_block, sizeof(int) * c->scan_cap));
  }
  if (!c->s26_cnt) {
    CUDA_TRY(cudaMalloc(&c->s26_cnt, 32 * sizeof(int)));
    CUDA_TRY(cudaMallocHost(&c->s26_hcnt, 32 * sizeof(int)));
  }
  k_s26_count<<<ntiles, S26_T, 0, st>>>(c->xq, n, G, c->scan_block, ntiles);
  k_s26_offsets<<<27, S26_T, 0, st>>>(c->scan_block, ntiles, c->s26_cnt);
  CUDA_TRY(cudaMemcpyAsync(c->s26_hcnt, c->s26_cnt, 27 * sizeof(int), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  S26Base first;
  int tot = 0;
  for (int d = 0; d < 27; d++) { first.v[d] = n + tot; tot += (d == 13) ? 0 : c->s26_hcnt[d]; }
  if (n + tot + 1 > c->nmax) TRY(neigh_grow_atoms(c, n + tot + 1));
  if (tot)
    k_s26_fill<<<ntiles, S26_T, 0, st>>>(c->xq, c->tag, n, G, c->scan_block, ntiles, first, c->ghost_src, c->ghost_pbc, c->prd[0],
                                         c->prd[1], c->prd[2]);
  c->launches += 3;
  CUDA_TRY(cudaGetLastError());
  c->nghost = tot;
  return B200MD_OK;
}

int comm_borders(b200md_ctx *c) {
  if (c->h26.on) return h26_borders(c);
  if (self26_usable(c)) return comm_borders_self26(c);
  cudaStream_t st = c->stream;
  c->nghost = 0;
  c->all_self = true;
  for (auto &s : c->swaps) if (s.sendproc != c->rank || s.recvproc != c->rank) c->all_self = false;
  int iswap = 0;
  for (int dim = 0; dim < 3; dim++) {
    int nlast = 0, nfirst = 0;
    const int twoneed = 2 * c->sched.maxneed[dim];
    for (int ineed = 0; ineed < twoneed; ineed += 2, iswap += 2) {
      Swap *sw[2] = {&c->swaps[iswap], &c->swaps[iswap + 1]};
      nfirst = nlast; nlast = c->nlocal + c->nghost;
      const int n = nlast - nfirst;
      int nsend[2] = {0, 0};
      SlabPair P;
      for (int k = 0; k < 2; k++) { P.lo[k] = sw[k]->slablo; P.hi[k] = sw[k]->slabhi; P.use[k] = sw[k]->sendflag ? 1 : 0; }
      const int ntiles = (n + SLAB_TILE - 1) / SLAB_TILE;
      if ((P.use[0] || P.use[1]) && n > 0) {
        if (2 * ntiles + 2 > c->scan_cap) {
          if (c->scan_block) CUDA_TRY(cudaFree(c->scan_block));
          c->scan_cap = 2 * ntiles + 64;
          CUDA_TRY(cudaMalloc(&c->scan_block, sizeof(int) * c->scan_cap));
        }
        int *tile_cnt = c->scan_block, *totals = c->d_flags + 6;
        k_slab_count2<<<ntiles, SLAB_T, 0, st>>>(c->xq, nfirst, n, sw[0]->dim, P, tile_cnt, ntiles);
        k_slab_offsets<<<1, SLAB_T, 0, st>>>(tile_cnt, ntiles, totals);
        CUDA_TRY(cudaMemcpyAsync(c->h_flags + 6, totals, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
        CUDA_TRY(cudaStreamSynchronize(st));
        nsend[0] = c->h_flags[6]; nsend[1] = c->h_flags[7];
        TRY(grow_sendlist(*sw[0], nsend[0]));
        TRY(grow_sendlist(*sw[1], nsend[1]));
        if (nsend[0] + nsend[1])
          k_slab_compact2<<<ntiles, SLAB_T, 0, st>>>(c->xq, nfirst, n, sw[0]->dim, P, tile_cnt, ntiles, sw[0]->sendlist,
                                                    sw[1]->sendlist);
        c->launches += 3;
      }
      // exchange per swap, in swap order, exactly like CommBrick::borders (comm_brick.cpp:917-960)
      for (int k = 0; k < 2; k++) {
        Swap &s = *sw[k];
        const int ns = nsend[k];
        s.sendnum = ns;
        int nrecv = ns;
        const bool self = (s.sendproc == c->rank);
        if (!self) {
          NcclState *nc = c->nccl;
          if (!nc) { b200md_set_error("b200md: comm not initialised (b200md_comm_init)"); return B200MD_ERR; }
          nc->h_counts[0] = ns;
          CUDA_TRY(cudaMemcpyAsync(nc->d_counts, nc->h_counts, sizeof(int), cudaMemcpyHostToDevice, st));
          NCCL_TRY(ncclGroupStart());
          NCCL_TRY(ncclSend(nc->d_counts, 1, ncclInt, s.sendproc, nc->comm, st));
          NCCL_TRY(ncclRecv(nc->d_counts + 1, 1, ncclInt, s.recvproc, nc->comm, st));
          NCCL_TRY(ncclGroupEnd());
          CUDA_TRY(cudaMemcpyAsync(nc->h_counts + 1, nc->d_counts + 1, sizeof(int), cudaMemcpyDeviceToHost, st));
          CUDA_TRY(cudaStreamSynchronize(st));
          nrecv = nc->h_counts[1];
        }
        const int first = c->nlocal + c->nghost;
        if (first + nrecv > c->nmax) TRY(neigh_grow_atoms(c, first + nrecv));
        if (self) {
          if (ns)
            k_self_border<<<nblk(ns), 256, 0, st>>>(ns, s.sendlist, first, c->nlocal, c->xq, c->tag, c->ghost_src,
                                                   c->ghost_pbc, s.pbc[0], s.pbc[1], s.pbc[2], c->prd[0], c->prd[1], c->prd[2],
                                                   c->direct.on ? c->ghost_rank : nullptr, c->rank);
          c->launches++;
        } else {
          NcclState *nc = c->nccl;
          const bool owner = c->direct.on;
          TRY(ensure_buf(c->sendbuf, c->sendcap, (size_t)ns * 7 + 16));
          double4 *buf = reinterpret_cast<double4 *>(c->sendbuf);
          int *tagbuf = reinterpret_cast<int *>(c->sendbuf + (size_t)ns * 4);
          int *o_src = tagbuf + ns, *o_code = o_src + ns, *o_rank = o_code + ns;
          if (ns) {
            k_pack_border<<<nblk(ns), 256, 0, st>>>(ns, s.sendlist, c->xq, c->tag, buf, tagbuf, s.pbc[0], s.pbc[1],
                                                   s.pbc[2], c->prd[0], c->prd[1], c->prd[2]);
            if (owner)
              k_pack_border_owner<<<nblk(ns), 256, 0, st>>>(ns, s.sendlist, c->nlocal, c->rank, c->ghost_src, c->ghost_pbc,
                                                           c->ghost_rank, s.pbc[0], s.pbc[1], s.pbc[2], o_src, o_code, o_rank);
          }
          c->launches++;
          NCCL_TRY(ncclGroupStart());
          if (ns) {
            NCCL_TRY(ncclSend(buf, (size_t)ns * 4, ncclDouble, s.sendproc, nc->comm, st));
            NCCL_TRY(ncclSend(tagbuf, ns, ncclInt, s.sendproc, nc->comm, st));
            if (owner) {
              NCCL_TRY(ncclSend(o_src, ns, ncclInt, s.sendproc, nc->comm, st));
              NCCL_TRY(ncclSend(o_code, ns, ncclInt, s.sendproc, nc->comm, st));
              NCCL_TRY(ncclSend(o_rank, ns, ncclInt, s.sendproc, nc->comm, st));
            }
          }
          if (nrecv) {
            NCCL_TRY(ncclRecv(c->xq + first, (size_t)nrecv * 4, ncclDouble, s.recvproc, nc->comm, st));
            NCCL_TRY(ncclRecv(c->tag + first, nrecv, ncclInt, s.recvproc, nc->comm, st));
            if (owner) {
              NCCL_TRY(ncclRecv(c->ghost_src + first, nrecv, ncclInt, s.recvproc, nc->comm, st));
              NCCL_TRY(ncclRecv(c->ghost_pbc + first, nrecv, ncclInt, s.recvproc, nc->comm, st));
              NCCL_TRY(ncclRecv(c->ghost_rank + first, nrecv, ncclInt, s.recvproc, nc->comm, st));
            }
          }
          NCCL_TRY(ncclGroupEnd());
        }
        s.recvnum = nrecv;
        s.firstrecv = first;
        c->nghost += nrecv;
      }
    }
  }
  CUDA_TRY(cudaGetLastError());
  c->direct.ready = false;
  if (c->direct.on && !c->all_self) TRY(direct_setup(c));
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// forward comm of positions (every non-rebuild step) and of a per-atom scalar (EAM fp)
// ---------------------------------------------------------------------------------------
template <bool SCALAR>
static int forward_staged(b200md_ctx *c, double *arr) {
  cudaStream_t st = c->stream;
  NcclState *nc = c->nccl;
  const int rec = SCALAR ? 1 : 4;
  size_t off[2] = {0, 0};
  int iswap = 0;
  for (int dim = 0; dim < 3; dim++) {
    const int twoneed = 2 * c->sched.maxneed[dim];
    for (int pair = 0; pair < twoneed; pair += 2) {
      // both directions of the pair read only atoms older than this pair's ghosts
      size_t need = 0;
      for (int k = 0; k < 2; k++) { off[k] = need; need += (size_t)c->swaps[iswap + k].sendnum * rec; }
      TRY(ensure_buf(c->sendbuf, c->sendcap, need + 8));
      bool any_remote = false;
      for (int k = 0; k < 2; k++) {
        Swap &s = c->swaps[iswap + k];
        if (s.sendproc == c->rank) continue;
        any_remote = true;
        if (!s.sendnum) continue;
        if (SCALAR) k_scalar_pack<<<nblk(s.sendnum), 256, 0, st>>>(s.sendnum, s.sendlist, arr, c->sendbuf + off[k]);
        else k_pack_border<<<nblk(s.sendnum), 256, 0, st>>>(s.sendnum, s.sendlist, c->xq, nullptr,
                                                            reinterpret_cast<double4 *>(c->sendbuf + off[k]), nullptr,
                                                            s.pbc[0], s.pbc[1], s.pbc[2], c->prd[0], c->prd[1], c->prd[2]);
        c->launches++;
      }
      if (any_remote) {
        if (!nc) { b200md_set_error("b200md: comm not initialised (b200md_comm_init)"); return B200MD_ERR; }
        NCCL_TRY(ncclGroupStart());
        for (int k = 0; k < 2; k++) {
          Swap &s = c->swaps[iswap + k];
          if (s.sendproc == c->rank) continue;
          if (s.sendnum) NCCL_TRY(ncclSend(c->sendbuf + off[k], (size_t)s.sendnum * rec, ncclDouble, s.sendproc, nc->comm, st));
          if (s.recvnum) {
            void *dst = SCALAR ? (void *)(arr + s.firstrecv) : (void *)(c->xq + s.firstrecv);
            NCCL_TRY(ncclRecv(dst, (size_t)s.recvnum * rec, ncclDouble, s.recvproc, nc->comm, st));
          }
        }
        NCCL_TRY(ncclGroupEnd());
      }
      for (int k = 0; k < 2; k++) {
        Swap &s = c->swaps[iswap + k];
        if (s.sendproc != c->rank || !s.sendnum) continue;
        if (SCALAR) k_scalar_self<<<nblk(s.sendnum), 256, 0, st>>>(s.sendnum, s.sendlist, s.firstrecv, arr);
        else k_self_forward<<<nblk(s.sendnum), 256, 0, st>>>(s.sendnum, s.sendlist, s.firstrecv, c->xq, s.pbc[0], s.pbc[1],
                                                             s.pbc[2], c->prd[0], c->prd[1], c->prd[2]);
        c->launches++;
      }
      iswap += 2;
    }
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

int comm_forward(b200md_ctx *c) {
  if (c->nghost == 0 && c->all_self) return B200MD_OK;
  if (c->all_self) {
    k_forward_flat<<<nblk(c->nghost), 256, 0, c->stream>>>(c->nghost, c->nlocal, c->xq, c->ghost_src, c->ghost_pbc,
                                                          c->prd[0], c->prd[1], c->prd[2]);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    return B200MD_OK;
  }
  if (c->h26.ready) return h26_forward(c, nullptr);
  if (c->direct.ready) return direct_forward<false>(c, nullptr);
  return forward_staged<false>(c, nullptr);
}

int comm_forward_scalar(b200md_ctx *c, double *arr) {
  if (c->nghost == 0 && c->all_self) return B200MD_OK;
  if (c->all_self) {
    k_scalar_flat<<<nblk(c->nghost), 256, 0, c->stream>>>(c->nghost, c->nlocal, arr, c->ghost_src);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    return B200MD_OK;
  }
  if (c->h26.ready) return h26_forward(c, arr);
  if (c->direct.ready) return direct_forward<true>(c, arr);
  return forward_staged<true>(c, arr);
}

// ---------------------------------------------------------------------------------------
// exchange: atom migration between neighbouring sub-boxes (comm_brick.cpp:652-759)
// record = 8 doubles {x,y,z,type-bits, vx,vy,vz, (tag | image<<32)}
// ---------------------------------------------------------------------------------------
__global__ void k_exch_flag(const double4 *__restrict__ xq, int n, int dim, double lo, double hi, int *__restrict__ flag) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n) return;
  if (k == n) { flag[k] = 0; return; }
  const double x = comp(xq[k], dim);
  flag[k] = (x < lo || x >= hi) ? 1 : 0;  // comm_brick.cpp:706
}
__global__ void k_exch_split(int n, int nmax, const int *__restrict__ flag, const int *__restrict__ pos,
                             const double4 *__restrict__ xq, const double *__restrict__ v, const int *__restrict__ tag,
                             const int *__restrict__ image, double4 *__restrict__ xq_o, double *__restrict__ v_o,
                             int *__restrict__ tag_o, int *__restrict__ image_o, double *__restrict__ buf) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const double4 q = xq[k];
  const double vx = v[k], vy = v[k + nmax], vz = v[k + 2 * (size_t)nmax];
  if (flag[k]) {
    double *r = buf + (size_t)pos[k] * 8;
    r[0] = q.x; r[1] = q.y; r[2] = q.z; r[3] = q.w; r[4] = vx; r[5] = vy; r[6] = vz;
    r[7] = __longlong_as_double(((long long)(unsigned)tag[k]) | ((long long)image[k] << 32));
  } else {
    const int d = k - pos[k];
    xq_o[d] = q; v_o[d] = vx; v_o[d + nmax] = vy; v_o[d + 2 * (size_t)nmax] = vz;
    tag_o[d] = tag[k]; image_o[d] = image[k];
  }
}
__global__ void k_exch_recv_flag(const double *__restrict__ buf, int n, int dim, double lo, double hi, int *__restrict__ flag) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k > n) return;
  if (k == n) { flag[k] = 0; return; }
  const double x = buf[(size_t)k * 8 + dim];
  flag[k] = (x >= lo && x < hi) ? 1 : 0;  // comm_brick.cpp:745-752
}
__global__ void k_exch_unpack(const double *__restrict__ buf, int n, const int *__restrict__ flag, const int *__restrict__ pos,
                              int first, int nmax, double4 *__restrict__ xq, double *__restrict__ v, int *__restrict__ tag,
                              int *__restrict__ image) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n || !flag[k]) return;
  const double *r = buf + (size_t)k * 8;
  const int d = first + pos[k];
  xq[d] = make_double4(r[0], r[1], r[2], r[3]);
  v[d] = r[4]; v[d + nmax] = r[5]; v[d + 2 * (size_t)nmax] = r[6];
  const long long ti = __double_as_longlong(r[7]);
  tag[d] = (int)(ti & 0xffffffffll);
  image[d] = (int)(ti >> 32);
}

int comm_exchange(b200md_ctx *c) {
  c->nghost = 0;  // ghosts are discarded (comm_brick.cpp:672-676)
  if (c->nranks == 1) return B200MD_OK;
  if (c->h26.on) return h26_exchange(c);
  cudaStream_t st = c->stream;
  NcclState *nc = c->nccl;
  if (!nc) { b200md_set_error("b200md: comm not initialised (b200md_comm_init)"); return B200MD_ERR; }
  for (int dim = 0; dim < 3; dim++) {
    if (c->procgrid[dim] == 1) continue;
    const int n = c->nlocal;
    if (n + 1 > c->nmax) TRY(neigh_grow_atoms(c, n + 1));
    int *flag = c->perm, *pos = c->atom_bin;
    k_exch_flag<<<nblk(n + 1), 256, 0, st>>>(c->xq, n, dim, c->sublo[dim], c->subhi[dim], flag);
    TRY(dev_exclusive_scan(c, flag, pos, n + 1));
    CUDA_TRY(cudaMemcpyAsync(c->h_flags + 4, pos + n, sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    const int nsend = c->h_flags[4];
    TRY(ensure_buf(c->sendbuf, c->sendcap, (size_t)nsend * 8 + 8));
    if (n) {
      k_exch_split<<<nblk(n), 256, 0, st>>>(n, c->nmax, flag, pos, c->xq, c->v, c->tag, c->image, c->xq_tmp, c->v_tmp,
                                           c->tag_tmp, c->image_tmp, c->sendbuf);
      std::swap(c->xq, c->xq_tmp); std::swap(c->v, c->v_tmp);
      std::swap(c->tag, c->tag_tmp); std::swap(c->image, c->image_tmp);
    }
    c->launches += 2;
    c->nlocal = n - nsend;
    // counts: to the left neighbour always, to the right one when procgrid > 2
    const int left = c->procneigh[dim][0], right = c->procneigh[dim][1];
    const bool both = c->procgrid[dim] > 2;
    nc->h_counts[0] = nsend;
    CUDA_TRY(cudaMemcpyAsync(nc->d_counts, nc->h_counts, sizeof(int), cudaMemcpyHostToDevice, st));
    NCCL_TRY(ncclGroupStart());
    NCCL_TRY(ncclSend(nc->d_counts, 1, ncclInt, left, nc->comm, st));
    NCCL_TRY(ncclRecv(nc->d_counts + 1, 1, ncclInt, right, nc->comm, st));
    if (both) {
      NCCL_TRY(ncclSend(nc->d_counts, 1, ncclInt, right, nc->comm, st));
      NCCL_TRY(ncclRecv(nc->d_counts + 2, 1, ncclInt, left, nc->comm, st));
    }
    NCCL_TRY(ncclGroupEnd());
    CUDA_TRY(cudaMemcpyAsync(nc->h_counts + 1, nc->d_counts + 1, 2 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    const int nrecv1 = nc->h_counts[1], nrecv2 = both ? nc->h_counts[2] : 0;
    const int nrecv = nrecv1 + nrecv2;
    TRY(ensure_buf(c->recvbuf, c->recvcap, (size_t)nrecv * 8 + 8));
    NCCL_TRY(ncclGroupStart());
    if (nsend) NCCL_TRY(ncclSend(c->sendbuf, (size_t)nsend * 8, ncclDouble, left, nc->comm, st));
    if (nrecv1) NCCL_TRY(ncclRecv(c->recvbuf, (size_t)nrecv1 * 8, ncclDouble, right, nc->comm, st));
    if (both) {
      if (nsend) NCCL_TRY(ncclSend(c->sendbuf, (size_t)nsend * 8, ncclDouble, right, nc->comm, st));
      if (nrecv2) NCCL_TRY(ncclRecv(c->recvbuf + (size_t)nrecv1 * 8, (size_t)nrecv2 * 8, ncclDouble, left, nc->comm, st));
    }
    NCCL_TRY(ncclGroupEnd());
    if (nrecv) {
      if (c->nlocal + nrecv + 1 > c->nmax) TRY(neigh_grow_atoms(c, c->nlocal + nrecv + 1));
      flag = c->perm; pos = c->atom_bin;
      k_exch_recv_flag<<<nblk(nrecv + 1), 256, 0, st>>>(c->recvbuf, nrecv, dim, c->sublo[dim], c->subhi[dim], flag);
      TRY(dev_exclusive_scan(c, flag, pos, nrecv + 1));
      k_exch_unpack<<<nblk(nrecv), 256, 0, st>>>(c->recvbuf, nrecv, flag, pos, c->nlocal, c->nmax, c->xq, c->v, c->tag, c->image);
      CUDA_TRY(cudaMemcpyAsync(c->h_flags + 4, pos + nrecv, sizeof(int), cudaMemcpyDeviceToHost, st));
      CUDA_TRY(cudaStreamSynchronize(st));
      c->nlocal += c->h_flags[4];
      c->launches += 2;
    }
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// reverse comm (half lists with newton on): ghost contributions summed back to owners
// ---------------------------------------------------------------------------------------
__global__ void k_rev_self3(int n, const int *__restrict__ list, int first, int nmax, double *__restrict__ f) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int j = list[k];
#pragma unroll
  for (int d = 0; d < 3; d++) atomicAdd(&f[j + d * (size_t)nmax], f[first + k + d * (size_t)nmax]);
}
__global__ void k_rev_self1(int n, const int *__restrict__ list, int first, double *__restrict__ a) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  atomicAdd(&a[list[k]], a[first + k]);
}
__global__ void k_rev_pack3(int n, int first, int nmax, const double *__restrict__ f, double *__restrict__ buf) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
#pragma unroll
  for (int d = 0; d < 3; d++) buf[(size_t)k * 3 + d] = f[first + k + d * (size_t)nmax];
}
__global__ void k_rev_unpack3(int n, const int *__restrict__ list, int nmax, double *__restrict__ f, const double *__restrict__ buf) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int j = list[k];
#pragma unroll
  for (int d = 0; d < 3; d++) atomicAdd(&f[j + d * (size_t)nmax], buf[(size_t)k * 3 + d]);
}
__global__ void k_rev_unpack1(int n, const int *__restrict__ list, double *__restrict__ a, const double *__restrict__ buf) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  atomicAdd(&a[list[k]], buf[k]);
}

// swaps in reverse order, one at a time; later (earlier-numbered) swaps see the result of this
// one exactly like comm_brick.cpp:610-638.  The adds are FP64 RED because, with compute/halo
// overlap, interior tiles of the force kernel accumulate into the same owned atoms concurrently.
template <int NV>
static int reverse_staged(b200md_ctx *c, double *arr) {
  cudaStream_t st = c->stream;
  NcclState *nc = c->nccl;
  for (int iswap = (int)c->swaps.size() - 1; iswap >= 0; iswap--) {
    Swap &s = c->swaps[iswap];
    if (s.sendproc == c->rank) {
      if (!s.sendnum) continue;
      if (NV == 3) k_rev_self3<<<nblk(s.sendnum), 256, 0, st>>>(s.sendnum, s.sendlist, s.firstrecv, c->nmax, arr);
      else k_rev_self1<<<nblk(s.sendnum), 256, 0, st>>>(s.sendnum, s.sendlist, s.firstrecv, arr);
      c->launches++;
      continue;
    }
    if (!nc) { b200md_set_error("b200md: comm not initialised (b200md_comm_init)"); return B200MD_ERR; }
    TRY(ensure_buf(c->sendbuf, c->sendcap, (size_t)s.recvnum * NV + 8));
    TRY(ensure_buf(c->recvbuf, c->recvcap, (size_t)s.sendnum * NV + 8));
    const double *src = arr + s.firstrecv;
    if (NV == 3 && s.recvnum) { k_rev_pack3<<<nblk(s.recvnum), 256, 0, st>>>(s.recvnum, s.firstrecv, c->nmax, arr, c->sendbuf); src = c->sendbuf; c->launches++; }
    NCCL_TRY(ncclGroupStart());
    if (s.recvnum) NCCL_TRY(ncclSend(src, (size_t)s.recvnum * NV, ncclDouble, s.recvproc, nc->comm, st));
    if (s.sendnum) NCCL_TRY(ncclRecv(c->recvbuf, (size_t)s.sendnum * NV, ncclDouble, s.sendproc, nc->comm, st));
    NCCL_TRY(ncclGroupEnd());
    if (s.sendnum) {
      if (NV == 3) k_rev_unpack3<<<nblk(s.sendnum), 256, 0, st>>>(s.sendnum, s.sendlist, c->nmax, arr, c->recvbuf);
      else k_rev_unpack1<<<nblk(s.sendnum), 256, 0, st>>>(s.sendnum, s.sendlist, arr, c->recvbuf);
      c->launches++;
    }
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

// single rank: every ghost adds straight into its flattened owner (ghost_src), one launch instead
// of one per swap; same sum as the staged order up to FP64 RED ordering
template <int NV>
__global__ void k_reverse_flat(int nghost, int nlocal, int nmax, double *__restrict__ a, const int *__restrict__ ghost_src) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nghost) return;
  const int o = ghost_src[nlocal + g];
#pragma unroll
  for (int d = 0; d < NV; d++) atomicAdd(&a[o + d * (size_t)nmax], a[nlocal + g + d * (size_t)nmax]);
}

int comm_reverse_force(b200md_ctx *c) {
  // newton_pair off: nothing on ghosts — unless bonds (newton_bond on) put their partner's share there,
  // which on one rank never happens (the tag map prefers the owned copy, bond.cu)
  if (c->list_kind == B200MD_LIST_FULL && !(c->topo.molecular && c->topo.fene && !c->all_self)) return B200MD_OK;
  if (c->all_self) {
    if (c->nghost == 0) return B200MD_OK;
    k_reverse_flat<3><<<nblk(c->nghost), 256, 0, c->stream>>>(c->nghost, c->nlocal, c->nmax, c->f, c->ghost_src);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    return B200MD_OK;
  }
  if (c->h26.ready) return h26_reverse3(c, c->f);
  if (c->direct.ready) return direct_reverse<3>(c, c->f);
  return reverse_staged<3>(c, c->f);
}
int comm_reverse_scalar(b200md_ctx *c, double *arr) {
  if (c->list_kind == B200MD_LIST_FULL) return B200MD_OK;
  return comm_reverse_column(c, arr);
}
// ghost rows of one per-atom column summed into their owners, whatever the list kind
int comm_reverse_column(b200md_ctx *c, double *arr) {
  if (c->all_self) {
    if (c->nghost == 0) return B200MD_OK;
    k_reverse_flat<1><<<nblk(c->nghost), 256, 0, c->stream>>>(c->nghost, c->nlocal, c->nmax, arr, c->ghost_src);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    return B200MD_OK;
  }
  if (c->h26.ready) return h26_reverse1(c, arr);
  if (c->direct.ready) return direct_reverse<1>(c, arr);
  return reverse_staged<1>(c, arr);
}
int comm_reverse_rho(b200md_ctx *c) { return comm_reverse_scalar(c, c->rho); }
