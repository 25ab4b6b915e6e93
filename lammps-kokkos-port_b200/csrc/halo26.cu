// halo26.cu — the "direct" halo on several ranks: ghosts and migrants go owner <-> holder in one hop.
//
// Reference behaviour reproduced (/root/reference/src):
//   CommBrick::exchange      comm_brick.cpp:652-759   atoms that left the sub-box move to the rank that now owns them
//   CommBrick::borders       comm_brick.cpp:771-979   ghost shell of width cutghost around the sub-box,
//   AtomVec::pack_border     atom_vec.cpp:794-830     ghost position = x + pbc*prd (one addition per crossed image)
//   CommBrick::forward_comm  comm_brick.cpp:538-591   per-step refresh of the ghost positions (and of EAM's fp)
//   CommBrick::reverse_comm  comm_brick.cpp:598-639   ghost forces (and EAM's rho) summed into their owners
// The reference forwards ghosts dimension by dimension (x, then y including the x-ghosts, then z): corner and edge
// ghosts reach their holder through up to three dependent messages, and every rebuild pays six count + six payload
// round trips.  The slab tests of the three stages are independent per dimension (a forwarded ghost is tested with
// its unshifted coordinate of the next dimension against the same bounds its owner would use), so the staged ghost
// set is exactly the product set: an owned atom is a ghost in direction (dx,dy,dz), dx/dy/dz in {-1,0,+1} not all 0,
// iff for every non-zero component it lies in that dimension's send slab.  Here each owned atom is classified ONCE
// by those up to 26 directions; a direction maps to one peer rank and one periodic shift, the copies go straight to
// that rank, and the holder's ghost block of a peer is laid out exactly like the owner's send list, so per-step
// messages land in place (no unpack kernel).  Per rebuild: a header round (counts) and a payload round for the
// migration, the same for the ghosts — 4 NCCL groups and 4 host syncs instead of ~20 of each; per step: one pack
// kernel and ONE group with a message per peer rank.  Ghost positions are bit-identical to the staged path (the same
// single addition of prd per crossed image and dimension); tests/test_multirank_gloo.py restates the protocol on the
// CPU and scripts/mgpu_parity.py / bench.py's parity_check compare thermo with the reference on 2-8 GPUs.
// Requires maxneed == 1 in every dimension (sub-box wider than the ghost cutoff); otherwise the staged path runs.
#include "comm.h"

#define H26_CENTER 13
#define MIG_REC 8   // doubles per migrant: x y z type-bits vx vy vz (tag | image << 32)

struct H26Geom {
  double sublo[3], subhi[3];
  double left_lo[3], left_hi[3], right_lo[3], right_hi[3];   // sub-box intervals of the two neighbours per dimension
  int can_left[3], can_right[3];
  double slabhi_left[3], slablo_right[3];                     // CommBrick::setup slab bounds (first swap of a dimension)
  int send_left[3], send_right[3];
};
struct H26Table { int v[H26_NDIR + 1]; };
struct H26Shift { int s[H26_NDIR][3]; };

static inline int dir_index(int dx, int dy, int dz) { return (dx + 1) + 3 * (dy + 1) + 9 * (dz + 1); }
static inline int roundup8(int n) { return (n + 7) & ~7; }

void h26_free(b200md_ctx *c) {
  Halo26 &H = c->h26;
  void *dev[] = {H.d_cnt, H.d_hdr_send, H.d_hdr_recv, H.list, H.dead, H.send_idx, H.send_dir, H.sbuf, H.rbuf};
  for (void *p : dev) if (p) cudaFree(p);
  if (H.h_cnt) cudaFreeHost(H.h_cnt);
  if (H.h_hdr) cudaFreeHost(H.h_hdr);
  const bool on = H.on;
  H = Halo26();
  H.on = on;
}

// direction tables from the swap schedule (comm_setup); decides whether the direct halo can run
int h26_setup(b200md_ctx *c, bool want) {
  Halo26 &H = c->h26;
  H.ready = false;
  H.on = false;
  if (!want || c->nranks == 1) return B200MD_OK;
  int usable = 0;
  double shl[3], slr[3];
  TRY(b200md_halo_directions(c->boxlo, c->boxhi, c->periodic, c->procgrid, c->myloc, c->cutghost, H.peer, H.shift, shl, slr, &usable));
  if (!usable) return B200MD_OK;   // several ghost layers per side: staged path
  const int me = c->rank;
  H.peers.clear(); H.dirs_self.clear();
  for (int d = 0; d < H26_NDIR; d++) {
    if (H.peer[d] < 0) continue;
    if (H.peer[d] == me) H.dirs_self.push_back(d);
    else if (std::find(H.peers.begin(), H.peers.end(), H.peer[d]) == H.peers.end()) H.peers.push_back(H.peer[d]);
  }
  std::sort(H.peers.begin(), H.peers.end());
  H.npeers = (int)H.peers.size();
  H.dirs_to.assign(H.npeers, {}); H.dirs_from.assign(H.npeers, {});
  for (int d = 0; d < H26_NDIR; d++) {
    if (d == H26_CENTER) continue;
    // outgoing direction d -> peer[d]; incoming direction d comes from the rank at myloc - d
    if (H.peer[d] >= 0 && H.peer[d] != me)
      H.dirs_to[std::find(H.peers.begin(), H.peers.end(), H.peer[d]) - H.peers.begin()].push_back(d);
    const int comp[3] = {d % 3 - 1, (d / 3) % 3 - 1, d / 9 - 1};
    int loc[3] = {c->myloc[0], c->myloc[1], c->myloc[2]};
    bool ok = true;
    for (int k = 0; k < 3 && ok; k++) {
      if (!comp[k]) continue;
      loc[k] -= comp[k];
      if (loc[k] < 0 || loc[k] >= c->procgrid[k]) {
        if (!c->periodic[k]) { ok = false; break; }
        loc[k] = loc[k] < 0 ? c->procgrid[k] - 1 : 0;
      }
    }
    if (!ok) continue;
    const int src = (loc[0] * c->procgrid[1] + loc[1]) * c->procgrid[2] + loc[2];
    if (src == me) continue;
    auto it = std::find(H.peers.begin(), H.peers.end(), src);
    if (it == H.peers.end()) { b200md_set_error("b200md: direct halo: asymmetric neighbour table"); return B200MD_ERR; }
    H.dirs_from[it - H.peers.begin()].push_back(d);
  }
  for (int p = 0; p < H.npeers; p++)
    if ((int)H.dirs_to[p].size() > H26_HDR || (int)H.dirs_from[p].size() > H26_HDR) {
      b200md_set_error("b200md: direct halo: too many directions per peer"); return B200MD_ERR;
    }
  if (!H.d_cnt) {
    CUDA_TRY(cudaMalloc(&H.d_cnt, 64 * sizeof(int)));
    CUDA_TRY(cudaMallocHost(&H.h_cnt, 64 * sizeof(int)));
  }
  if (H.d_hdr_send) { cudaFree(H.d_hdr_send); cudaFree(H.d_hdr_recv); cudaFreeHost(H.h_hdr); }
  const size_t hb = sizeof(int) * (size_t)std::max(1, H.npeers) * H26_HDR;
  CUDA_TRY(cudaMalloc(&H.d_hdr_send, hb));
  CUDA_TRY(cudaMalloc(&H.d_hdr_recv, hb));
  CUDA_TRY(cudaMallocHost(&H.h_hdr, 2 * hb));
  for (int d = 0; d < H26_NDIR; d++) { H.cap_b[d] = 0; H.cap_m[d] = 0; }
  H.send_off.assign(H.npeers, 0); H.send_cnt.assign(H.npeers, 0);
  H.recv_off.assign(H.npeers, 0); H.recv_cnt.assign(H.npeers, 0);
  H.on = true;
  return B200MD_OK;
}

static H26Geom make_geom(const b200md_ctx *c) {
  H26Geom G;
  for (int k = 0; k < 3; k++) {
    G.sublo[k] = c->sublo[k]; G.subhi[k] = c->subhi[k];
    const int P = c->procgrid[k];
    int loc[3] = {c->myloc[0], c->myloc[1], c->myloc[2]};
    double lo[3], hi[3];
    int l = c->myloc[k] - 1, r = c->myloc[k] + 1;
    G.can_left[k] = P > 1 && (l >= 0 || c->periodic[k]);
    G.can_right[k] = P > 1 && (r < P || c->periodic[k]);
    if (l < 0) l = P - 1;
    if (r >= P) r = 0;
    loc[k] = l; b200md_local_box(c->boxlo, c->boxhi, c->procgrid, loc, lo, hi); G.left_lo[k] = lo[k]; G.left_hi[k] = hi[k];
    loc[k] = r; b200md_local_box(c->boxlo, c->boxhi, c->procgrid, loc, lo, hi); G.right_lo[k] = lo[k]; G.right_hi[k] = hi[k];
    G.slabhi_left[k] = c->swaps[2 * k].slabhi; G.slablo_right[k] = c->swaps[2 * k + 1].slablo;
    G.send_left[k] = c->swaps[2 * k].sendflag; G.send_right[k] = c->swaps[2 * k + 1].sendflag;
  }
  return G;
}

__device__ __forceinline__ double comp3(const double4 &q, int k) { return k == 0 ? q.x : (k == 1 ? q.y : q.z); }

// ---------------------------------------------------------------------------------------
// migration
// ---------------------------------------------------------------------------------------
// every owned atom: the direction of the sub-box that now contains it (comm_brick.cpp:706: x < sublo || x >= subhi);
// leavers are packed straight into their direction's segment and marked dead (the next bin sort drops them)
__global__ void k_h26_mig_classify(int n, int nmax, const double4 *__restrict__ xq, const double *__restrict__ v,
                                   const int *__restrict__ tag, const int *__restrict__ image, H26Geom G, H26Table cap,
                                   H26Table lo, int *__restrict__ cnt, int *__restrict__ dead, double *__restrict__ buf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double4 q = xq[i];
  int s[3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const double x = comp3(q, k);
    s[k] = 0;
    if (x < G.sublo[k] || x >= G.subhi[k]) {
      if (G.can_left[k] && x >= G.left_lo[k] && x < G.left_hi[k]) s[k] = -1;
      else if (G.can_right[k] && x >= G.right_lo[k] && x < G.right_hi[k]) s[k] = 1;
      else if (G.can_left[k] || G.can_right[k]) atomicOr(&cnt[60], 1);   // moved further than the neighbouring sub-box
    }
  }
  const int d = (s[0] + 1) + 3 * (s[1] + 1) + 9 * (s[2] + 1);
  if (d == H26_CENTER) { dead[i] = 0; return; }
  dead[i] = 1;
  const int slot = atomicAdd(&cnt[32 + d], 1);
  if (slot >= cap.v[d]) return;
  double *r = buf + (size_t)(lo.v[d] + slot) * MIG_REC;
  r[0] = q.x; r[1] = q.y; r[2] = q.z; r[3] = q.w;
  r[4] = v[i]; r[5] = v[i + nmax]; r[6] = v[i + 2 * (size_t)nmax];
  r[7] = __longlong_as_double(((long long)(unsigned)tag[i]) | ((long long)image[i] << 32));
}
__global__ void k_h26_mig_unpack(int n, const double *__restrict__ buf, int first, int nmax, double4 *__restrict__ xq,
                                 double *__restrict__ v, int *__restrict__ tag, int *__restrict__ image, int *__restrict__ dead) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const double *r = buf + (size_t)k * MIG_REC;
  const int d = first + k;
  xq[d] = make_double4(r[0], r[1], r[2], r[3]);
  v[d] = r[4]; v[d + nmax] = r[5]; v[d + 2 * (size_t)nmax] = r[6];
  const long long ti = __double_as_longlong(r[7]);
  tag[d] = (int)(ti & 0xffffffffll);
  image[d] = (int)(ti >> 32);
  dead[d] = 0;
}

static int ensure_dead(b200md_ctx *c) {
  Halo26 &H = c->h26;
  if (H.dead_cap >= c->nmax) return B200MD_OK;
  int *p = nullptr;
  CUDA_TRY(cudaMalloc(&p, sizeof(int) * (size_t)c->nmax));
  CUDA_TRY(cudaMemsetAsync(p, 0, sizeof(int) * (size_t)c->nmax, c->stream));
  if (H.dead) {
    CUDA_TRY(cudaMemcpyAsync(p, H.dead, sizeof(int) * (size_t)H.dead_cap, cudaMemcpyDeviceToDevice, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    CUDA_TRY(cudaFree(H.dead));
  }
  H.dead = p; H.dead_cap = c->nmax;
  return B200MD_OK;
}

// counts per direction to every peer, theirs back: one group of H26_HDR-int messages + one sync
static int exchange_headers(b200md_ctx *c, const int *counts, int *counts_in) {
  Halo26 &H = c->h26;
  NcclState *nc = c->nccl;
  cudaStream_t st = c->stream;
  int *hs = H.h_hdr, *hr = H.h_hdr + (size_t)H.npeers * H26_HDR;
  for (int p = 0; p < H.npeers; p++) {
    for (int k = 0; k < H26_HDR; k++) hs[p * H26_HDR + k] = 0;
    for (size_t k = 0; k < H.dirs_to[p].size(); k++) hs[p * H26_HDR + k] = counts[H.dirs_to[p][k]];
  }
  const size_t hb = sizeof(int) * (size_t)H.npeers * H26_HDR;
  CUDA_TRY(cudaMemcpyAsync(H.d_hdr_send, hs, hb, cudaMemcpyHostToDevice, st));
  NCCL_TRY(ncclGroupStart());
  for (int p = 0; p < H.npeers; p++) {
    NCCL_TRY(ncclSend(H.d_hdr_send + p * H26_HDR, H26_HDR, ncclInt, H.peers[p], nc->comm, st));
    NCCL_TRY(ncclRecv(H.d_hdr_recv + p * H26_HDR, H26_HDR, ncclInt, H.peers[p], nc->comm, st));
  }
  NCCL_TRY(ncclGroupEnd());
  CUDA_TRY(cudaMemcpyAsync(hr, H.d_hdr_recv, hb, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  for (int d = 0; d < H26_NDIR; d++) counts_in[d] = 0;
  for (int p = 0; p < H.npeers; p++)
    for (size_t k = 0; k < H.dirs_from[p].size(); k++) counts_in[H.dirs_from[p][k]] = hr[p * H26_HDR + k];
  return B200MD_OK;
}

int h26_exchange(b200md_ctx *c) {
  Halo26 &H = c->h26;
  NcclState *nc = c->nccl;
  if (!nc) { b200md_set_error("b200md: comm not initialised (b200md_comm_init)"); return B200MD_ERR; }
  cudaStream_t st = c->stream;
  c->nghost = 0;   // ghosts are discarded (comm_brick.cpp:672-676)
  H.ready = false;
  const int n = c->nlocal;
  TRY(ensure_dead(c));
  const H26Geom G = make_geom(c);
  int counts[H26_NDIR], counts_in[H26_NDIR];
  for (int attempt = 0;; attempt++) {
    // segments of the send buffer, in the order (peer, direction) the payload ops use
    H26Table cap, lo;
    int total = 0;
    for (int d = 0; d <= H26_NDIR; d++) { cap.v[d] = 0; lo.v[d] = 0; }
    for (int p = 0; p < H.npeers; p++)
      for (int d : H.dirs_to[p]) {
        if (H.cap_m[d] == 0) H.cap_m[d] = 2048;
        cap.v[d] = H.cap_m[d]; lo.v[d] = total; total += H.cap_m[d];
      }
    TRY(ensure_buf(c->sendbuf, c->sendcap, (size_t)total * MIG_REC + 8));
    CUDA_TRY(cudaMemsetAsync(H.d_cnt + 32, 0, 32 * sizeof(int), st));
    if (n) k_h26_mig_classify<<<(n + 255) / 256, 256, 0, st>>>(n, c->nmax, c->xq, c->v, c->tag, c->image, G, cap, lo, H.d_cnt,
                                                               H.dead, c->sendbuf);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(H.h_cnt + 32, H.d_cnt + 32, 32 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    if (H.h_cnt[60]) { b200md_set_error("Lost atoms: an atom moved further than the neighbouring sub-domain between two rebuilds"); return B200MD_ERR; }
    bool over = false;
    for (int d = 0; d < H26_NDIR; d++) {
      counts[d] = H.h_cnt[32 + d];
      if (counts[d] > H.cap_m[d]) {
        if (H.peer[d] < 0 || H.peer[d] == c->rank) { b200md_set_error("b200md: direct halo: migrant in an unused direction"); return B200MD_ERR; }
        H.cap_m[d] = roundup8(counts[d] * 2 + 1024); over = true;
      }
    }
    if (!over) {
      // the segment offsets of this attempt are what the payload ops read from
      TRY(exchange_headers(c, counts, counts_in));
      int nsend = 0, nrecv = 0;
      for (int d = 0; d < H26_NDIR; d++) { nsend += counts[d]; nrecv += counts_in[d]; }
      if (n + nrecv + 1 > c->nmax) { TRY(neigh_grow_atoms(c, n + nrecv + 1)); TRY(ensure_dead(c)); }
      TRY(ensure_buf(c->recvbuf, c->recvcap, (size_t)nrecv * MIG_REC + 8));
      NCCL_TRY(ncclGroupStart());
      size_t roff = 0;
      for (int p = 0; p < H.npeers; p++) {
        for (int d : H.dirs_to[p])
          if (counts[d]) NCCL_TRY(ncclSend(c->sendbuf + (size_t)lo.v[d] * MIG_REC, (size_t)counts[d] * MIG_REC, ncclDouble, H.peers[p], nc->comm, st));
        for (int d : H.dirs_from[p])
          if (counts_in[d]) { NCCL_TRY(ncclRecv(c->recvbuf + roff, (size_t)counts_in[d] * MIG_REC, ncclDouble, H.peers[p], nc->comm, st)); roff += (size_t)counts_in[d] * MIG_REC; }
      }
      NCCL_TRY(ncclGroupEnd());
      if (nrecv) {
        k_h26_mig_unpack<<<(nrecv + 255) / 256, 256, 0, st>>>(nrecv, c->recvbuf, n, c->nmax, c->xq, c->v, c->tag, c->image, H.dead);
        c->launches++;
      }
      CUDA_TRY(cudaGetLastError());
      c->nlocal = n + nrecv;   // leavers included until the sort drops them
      H.ndead = nsend;
      return B200MD_OK;
    }
    if (attempt > 4) { b200md_set_error("b200md: direct halo: migrant buffers keep overflowing"); return B200MD_ERR; }
    H.nredo++;
  }
}

// ---------------------------------------------------------------------------------------
// ghosts
// ---------------------------------------------------------------------------------------
// every owned atom: the directions it is a ghost in = product over the dimensions of {0, -1 if in the left send slab,
// +1 if in the right one} minus (0,0,0); slab test as comm_brick.cpp:838 (x >= lo && x <= hi, lo = -BIG / hi = BIG
// for the first swap of a dimension)
__global__ void k_h26_bor_classify(int n, const double4 *__restrict__ xq, H26Geom G, H26Table cap, H26Table lo,
                                   int *__restrict__ cnt, int *__restrict__ list) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double4 q = xq[i];
  bool L[3], R[3];
  bool any = false;
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const double x = comp3(q, k);
    L[k] = G.send_left[k] && x <= G.slabhi_left[k];
    R[k] = G.send_right[k] && x >= G.slablo_right[k];
    any |= L[k] | R[k];
  }
  if (!any) return;
  for (int dz = -1; dz <= 1; dz++) {
    if ((dz < 0 && !L[2]) || (dz > 0 && !R[2])) continue;
    for (int dy = -1; dy <= 1; dy++) {
      if ((dy < 0 && !L[1]) || (dy > 0 && !R[1])) continue;
      for (int dx = -1; dx <= 1; dx++) {
        if ((dx < 0 && !L[0]) || (dx > 0 && !R[0])) continue;
        const int d = (dx + 1) + 3 * (dy + 1) + 9 * (dz + 1);
        if (d == H26_CENTER) continue;
        const int slot = atomicAdd(&cnt[d], 1);
        if (slot < cap.v[d]) list[lo.v[d] + slot] = i;
      }
    }
  }
}
// capacity-spaced candidate lists -> the dense send list, ordered (remote peers ascending, direction), then the
// copies that stay on this rank.  seg.v[k] = first dense entry of the k-th segment, dirs.v[k] its direction.
__global__ void k_h26_compact(int nsend, int nseg, H26Table seg, H26Table dirs, H26Table lo, const int *__restrict__ list,
                              int *__restrict__ send_idx, int *__restrict__ send_dir) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nsend) return;
  int k = 0;
  while (k + 1 < nseg && e >= seg.v[k + 1]) k++;
  const int d = dirs.v[k];
  send_idx[e] = list[lo.v[d] + (e - seg.v[k])];
  send_dir[e] = d;
}
// positions (+ tags at a rebuild) of the outgoing copies with their direction's shift; remote copies into the send
// buffer, copies that stay here straight into their ghost slots
template <bool TAGS>
__global__ void k_h26_pack_x(int nsend, int nremote, int self_first, const int *__restrict__ send_idx,
                             const int *__restrict__ send_dir, H26Shift S, double prdx, double prdy, double prdz,
                             double4 *__restrict__ xq, int *__restrict__ tag, double4 *__restrict__ posbuf, int *__restrict__ tagbuf) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nsend) return;
  const int i = send_idx[e], d = send_dir[e];
  double4 q = xq[i];
  q.x = shift_n(q.x, S.s[d][0], prdx); q.y = shift_n(q.y, S.s[d][1], prdy); q.z = shift_n(q.z, S.s[d][2], prdz);
  if (e < nremote) { posbuf[e] = q; if (TAGS) tagbuf[e] = tag[i]; }
  else { xq[self_first + e - nremote] = q; if (TAGS) tag[self_first + e - nremote] = tag[i]; }
}
__global__ void k_h26_pack_s(int nsend, int nremote, int self_first, const int *__restrict__ send_idx, double *__restrict__ a,
                             double *__restrict__ buf) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nsend) return;
  const double val = a[send_idx[e]];
  if (e < nremote) buf[e] = val; else a[self_first + e - nremote] = val;
}
template <int NV>
__global__ void k_h26_rev_pack(int nremote, int first, int nmax, const double *__restrict__ a, double *__restrict__ buf) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= nremote) return;
#pragma unroll
  for (int k = 0; k < NV; k++) buf[(size_t)g * NV + k] = a[first + g + k * (size_t)nmax];
}
// FP64 RED: with compute/halo overlap the interior tiles of the force kernel add into the same owned atoms concurrently
template <int NV>
__global__ void k_h26_rev_unpack(int nsend, int nremote, int self_first, int nmax, const int *__restrict__ send_idx,
                                 double *__restrict__ a, const double *__restrict__ buf) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= nsend) return;
  const int i = send_idx[e];
#pragma unroll
  for (int k = 0; k < NV; k++) {
    const double val = e < nremote ? buf[(size_t)e * NV + k] : a[self_first + e - nremote + k * (size_t)nmax];
    atomicAdd(&a[i + k * (size_t)nmax], val);
  }
}

static H26Shift make_shift(const Halo26 &H) {
  H26Shift S;
  for (int d = 0; d < H26_NDIR; d++) for (int k = 0; k < 3; k++) S.s[d][k] = H.shift[d][k];
  return S;
}

int h26_borders(b200md_ctx *c) {
  Halo26 &H = c->h26;
  NcclState *nc = c->nccl;
  if (!nc) { b200md_set_error("b200md: comm not initialised (b200md_comm_init)"); return B200MD_ERR; }
  cudaStream_t st = c->stream;
  c->nghost = 0;
  c->all_self = false;
  H.ready = false;
  const int n = c->nlocal;
  const H26Geom G = make_geom(c);
  int counts[H26_NDIR], counts_in[H26_NDIR];
  H26Table cap, lo;
  for (int attempt = 0;; attempt++) {
    int total = 0;
    for (int d = 0; d <= H26_NDIR; d++) { cap.v[d] = 0; lo.v[d] = 0; }
    for (int d = 0; d < H26_NDIR; d++) {
      if (H.peer[d] < 0) continue;
      if (H.cap_b[d] == 0) {
        // first guess: the slab's share of the owned atoms (face: cutghost / width), generously
        double frac = 1.0;
        const int comp[3] = {d % 3 - 1, (d / 3) % 3 - 1, d / 9 - 1};
        for (int k = 0; k < 3; k++) if (comp[k]) frac *= std::min(1.0, c->cutghost / (c->subhi[k] - c->sublo[k]));
        H.cap_b[d] = roundup8((int)(n * frac * 1.5) + 1024);
      }
      cap.v[d] = H.cap_b[d]; lo.v[d] = total; total += H.cap_b[d];
    }
    if (total > H.list_cap) {
      if (H.list) CUDA_TRY(cudaFree(H.list));
      H.list_cap = total + total / 8;
      CUDA_TRY(cudaMalloc(&H.list, sizeof(int) * (size_t)H.list_cap));
    }
    CUDA_TRY(cudaMemsetAsync(H.d_cnt, 0, 32 * sizeof(int), st));
    if (n) k_h26_bor_classify<<<(n + 255) / 256, 256, 0, st>>>(n, c->xq, G, cap, lo, H.d_cnt, H.list);
    c->launches++;
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(H.h_cnt, H.d_cnt, 32 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    bool over = false;
    for (int d = 0; d < H26_NDIR; d++) {
      counts[d] = H.h_cnt[d];
      if (counts[d] > H.cap_b[d]) { H.cap_b[d] = roundup8(counts[d] + counts[d] / 4 + 1024); over = true; }
    }
    if (!over) break;
    if (attempt > 4) { b200md_set_error("b200md: direct halo: ghost lists keep overflowing"); return B200MD_ERR; }
    H.nredo++;
  }
  for (int d = 0; d < H26_NDIR; d++)   // next rebuild: room for a quarter more than this one needed
    if (H.peer[d] >= 0) H.cap_b[d] = std::max(H.cap_b[d], roundup8(counts[d] + counts[d] / 4 + 1024));
  TRY(exchange_headers(c, counts, counts_in));

  // dense send list: (remote peers ascending, their directions ascending), then the copies that stay here
  H26Table seg, dirs;
  int nseg = 0, e = 0;
  for (int p = 0; p < H.npeers; p++) {
    H.send_off[p] = e;
    for (int d : H.dirs_to[p]) { seg.v[nseg] = e; dirs.v[nseg] = d; nseg++; e += counts[d]; }
    H.send_cnt[p] = e - H.send_off[p];
  }
  H.nsend_remote = e;
  for (int d : H.dirs_self) { seg.v[nseg] = e; dirs.v[nseg] = d; nseg++; e += counts[d]; }
  H.nsend = e;
  seg.v[nseg] = e;
  int g = 0;
  for (int p = 0; p < H.npeers; p++) {
    H.recv_off[p] = g;
    for (int d : H.dirs_from[p]) g += counts_in[d];
    H.recv_cnt[p] = g - H.recv_off[p];
  }
  H.nghost_remote = g;
  const int nghost = g + (H.nsend - H.nsend_remote);
  if (n + nghost + 1 > c->nmax) TRY(neigh_grow_atoms(c, n + nghost + 1));
  if (H.nsend > H.send_cap) {
    if (H.send_idx) { CUDA_TRY(cudaFree(H.send_idx)); CUDA_TRY(cudaFree(H.send_dir)); }
    H.send_cap = H.nsend + H.nsend / 4 + 1024;
    CUDA_TRY(cudaMalloc(&H.send_idx, sizeof(int) * (size_t)H.send_cap));
    CUDA_TRY(cudaMalloc(&H.send_dir, sizeof(int) * (size_t)H.send_cap));
  }
  // positions as double4 records, tags behind them
  TRY(ensure_buf(c->sendbuf, c->sendcap, (size_t)H.nsend_remote * 5 + 16));
  double4 *posbuf = reinterpret_cast<double4 *>(c->sendbuf);
  int *tagbuf = reinterpret_cast<int *>(c->sendbuf + (size_t)H.nsend_remote * 4);
  if (H.nsend) {
    k_h26_compact<<<(H.nsend + 255) / 256, 256, 0, st>>>(H.nsend, nseg, seg, dirs, lo, H.list, H.send_idx, H.send_dir);
    k_h26_pack_x<true><<<(H.nsend + 255) / 256, 256, 0, st>>>(H.nsend, H.nsend_remote, n + H.nghost_remote, H.send_idx, H.send_dir,
                                                              make_shift(H), c->prd[0], c->prd[1], c->prd[2], c->xq, c->tag, posbuf, tagbuf);
    c->launches += 2;
  }
  NCCL_TRY(ncclGroupStart());
  for (int p = 0; p < H.npeers; p++) {
    if (H.send_cnt[p]) {
      NCCL_TRY(ncclSend(posbuf + H.send_off[p], (size_t)H.send_cnt[p] * 4, ncclDouble, H.peers[p], nc->comm, st));
      NCCL_TRY(ncclSend(tagbuf + H.send_off[p], H.send_cnt[p], ncclInt, H.peers[p], nc->comm, st));
    }
    if (H.recv_cnt[p]) {
      NCCL_TRY(ncclRecv(c->xq + n + H.recv_off[p], (size_t)H.recv_cnt[p] * 4, ncclDouble, H.peers[p], nc->comm, st));
      NCCL_TRY(ncclRecv(c->tag + n + H.recv_off[p], H.recv_cnt[p], ncclInt, H.peers[p], nc->comm, st));
    }
  }
  NCCL_TRY(ncclGroupEnd());
  CUDA_TRY(cudaGetLastError());
  c->nghost = nghost;
  H.ready = true;
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// per-step traffic: one pack kernel, one group (a message per peer rank), messages land in place
// ---------------------------------------------------------------------------------------
int h26_forward(b200md_ctx *c, double *scalar) {
  Halo26 &H = c->h26;
  NcclState *nc = c->nccl;
  cudaStream_t st = c->stream;
  const int n = c->nlocal, self_first = n + H.nghost_remote;
  if (!H.nsend && !H.nghost_remote) return B200MD_OK;
  TRY(ensure_buf(c->sendbuf, c->sendcap, (size_t)H.nsend_remote * 4 + 16));
  if (H.nsend) {
    if (scalar) k_h26_pack_s<<<(H.nsend + 255) / 256, 256, 0, st>>>(H.nsend, H.nsend_remote, self_first, H.send_idx, scalar, c->sendbuf);
    else k_h26_pack_x<false><<<(H.nsend + 255) / 256, 256, 0, st>>>(H.nsend, H.nsend_remote, self_first, H.send_idx, H.send_dir, make_shift(H),
                                                                    c->prd[0], c->prd[1], c->prd[2], c->xq, nullptr,
                                                                    reinterpret_cast<double4 *>(c->sendbuf), nullptr);
    c->launches++;
  }
  const int rec = scalar ? 1 : 4;
  NCCL_TRY(ncclGroupStart());
  for (int p = 0; p < H.npeers; p++) {
    if (H.send_cnt[p]) NCCL_TRY(ncclSend(c->sendbuf + (size_t)H.send_off[p] * rec, (size_t)H.send_cnt[p] * rec, ncclDouble, H.peers[p], nc->comm, st));
    if (H.recv_cnt[p]) {
      void *dst = scalar ? (void *)(scalar + n + H.recv_off[p]) : (void *)(c->xq + n + H.recv_off[p]);
      NCCL_TRY(ncclRecv(dst, (size_t)H.recv_cnt[p] * rec, ncclDouble, H.peers[p], nc->comm, st));
    }
  }
  NCCL_TRY(ncclGroupEnd());
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

template <int NV>
static int h26_reverse_t(b200md_ctx *c, double *arr) {
  Halo26 &H = c->h26;
  NcclState *nc = c->nccl;
  cudaStream_t st = c->stream;
  const int n = c->nlocal, self_first = n + H.nghost_remote;
  if (!H.nsend && !H.nghost_remote) return B200MD_OK;
  TRY(ensure_buf(c->sendbuf, c->sendcap, (size_t)H.nghost_remote * NV + 16));
  TRY(ensure_buf(c->recvbuf, c->recvcap, (size_t)H.nsend_remote * NV + 16));
  const double *src = arr + n;   // NV == 1: the ghost block of a peer is contiguous, sent as it lies
  if (NV > 1 && H.nghost_remote) {
    k_h26_rev_pack<NV><<<(H.nghost_remote + 255) / 256, 256, 0, st>>>(H.nghost_remote, n, c->nmax, arr, c->sendbuf);
    c->launches++;
    src = c->sendbuf;
  }
  NCCL_TRY(ncclGroupStart());
  for (int p = 0; p < H.npeers; p++) {
    if (H.recv_cnt[p]) NCCL_TRY(ncclSend(src + (size_t)H.recv_off[p] * NV, (size_t)H.recv_cnt[p] * NV, ncclDouble, H.peers[p], nc->comm, st));
    if (H.send_cnt[p]) NCCL_TRY(ncclRecv(c->recvbuf + (size_t)H.send_off[p] * NV, (size_t)H.send_cnt[p] * NV, ncclDouble, H.peers[p], nc->comm, st));
  }
  NCCL_TRY(ncclGroupEnd());
  if (H.nsend) {
    k_h26_rev_unpack<NV><<<(H.nsend + 255) / 256, 256, 0, st>>>(H.nsend, H.nsend_remote, self_first, c->nmax, H.send_idx, arr, c->recvbuf);
    c->launches++;
  }
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}
int h26_reverse3(b200md_ctx *c, double *arr) { return h26_reverse_t<3>(c, arr); }
int h26_reverse1(b200md_ctx *c, double *arr) { return h26_reverse_t<1>(c, arr); }
