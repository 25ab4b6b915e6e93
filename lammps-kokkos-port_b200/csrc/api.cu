// api.cu — the extern "C" entry points of include/b200md.h (context, atoms, neighbor
// schedule, thermo, whole-loop driver).  Kernels live in neigh.cu / comm.cu / pair_*.cu /
// nve.cu; host-only geometry in host.cpp.
//
// Step order follows Verlet::setup (/root/reference/src/verlet.cpp:93-162) and Verlet::run
// (:229-360); the rebuild decision follows Neighbor::decide/check_distance
// (src/neighbor.cpp:2309-2390).
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "ctx.h"
#include "util.cuh"

int vel_adjust(b200md_ctx *c, double sx, double sy, double sz, double factor);  // nve.cu
int pair_eam_config_impl(b200md_ctx *c, int ntypes, int nrho, int nr, double rdr, double rdrho, double rhomax,
                         double cutforcesq, int nfrho, const double *frho_spline, int nrhor,
                         const double *rhor_spline, int nz2r, const double *z2r_spline, const int *type2frho,
                         const int *type2rhor, const int *type2z2r, const double *scale);  // pair_eam.cu
void pair_eam_free(b200md_ctx *c);

static thread_local std::string g_error;
void b200md_set_error(const std::string &msg) { g_error = msg; }
extern "C" const char *b200md_last_error(void) { return g_error.c_str(); }
extern "C" const char *b200md_version(void) { return "b200md 0.2 (sm_100a)"; }

#define NEED(cond, msg) do { if (!(cond)) { b200md_set_error(msg); return B200MD_ERR; } } while (0)

extern "C" int b200md_device_count(int *count) {
  if (!count) { b200md_set_error("b200md: null argument"); return B200MD_ERR; }
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; }
  *count = n;
  return B200MD_OK;
}

extern "C" int b200md_create(int device, int rank, int nranks, b200md_ctx **out) {
  NEED(out, "b200md: null output pointer");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    b200md_set_error("b200md: no CUDA device (this library has no CPU fallback)");
    return B200MD_ERR;
  }
  NEED(device >= 0 && device < ndev, "b200md: device index out of range");
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major < 10) {
    b200md_set_error(std::string("b200md: device ") + prop.name + " is not sm_100 (B200); kernels are built for sm_100a only");
    return B200MD_ERR;
  }
  CUDA_TRY(cudaSetDevice(device));
  b200md_ctx *c = new b200md_ctx();
  c->device = device; c->rank = rank; c->nranks = nranks;
  memset(&c->timing, 0, sizeof(c->timing));
  memset(&c->bin, 0, sizeof(c->bin));
  memset(&c->sched, 0, sizeof(c->sched));
  CUDA_TRY(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  {  // halo traffic gets the highest priority so its small kernels slip in between force blocks
    int lo = 0, hi = 0;
    CUDA_TRY(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CUDA_TRY(cudaStreamCreateWithPriority(&c->comm_stream, cudaStreamNonBlocking, hi));
    CUDA_TRY(cudaStreamCreateWithPriority(&c->bnd_stream, cudaStreamNonBlocking, hi));   // boundary tiles: ahead of the interior kernel
  }
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_x, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_halo, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_fp, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_fphalo, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_bnd, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_clr, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_rev, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_rb, cudaEventDisableTiming));
  CUDA_TRY(cudaEventCreateWithFlags(&c->ev_rr, cudaEventDisableTiming));
  // runtime knobs are set through b200md_set_option (the library reads no environment variables);
  // per-step halo: owner <-> holder in one exchange ("direct", default on several ranks) or the
  // reference's three forwarding stages ("staged")
  c->halo_mode = 2;
  CUDA_TRY(cudaMalloc(&c->d_result, 16 * sizeof(double)));
  CUDA_TRY(cudaMallocHost(&c->h_result, 16 * sizeof(double)));
  CUDA_TRY(cudaMalloc(&c->d_flags, 8 * sizeof(int)));
  CUDA_TRY(cudaMemset(c->d_flags, 0, 8 * sizeof(int)));
  CUDA_TRY(cudaMallocHost(&c->h_flags, 8 * sizeof(int)));
  memset(c->h_flags, 0, 8 * sizeof(int));
  *out = c;
  return B200MD_OK;
}

extern "C" int b200md_destroy(b200md_ctx *c) {
  if (!c) return B200MD_OK;
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  if (c->copy_stream) {
    cudaStreamSynchronize(c->copy_stream);
    cudaStreamDestroy(c->copy_stream);
    cudaEventDestroy(c->ev_packed); cudaEventDestroy(c->ev_dl_x); cudaEventDestroy(c->ev_dl_all);
    for (cudaEvent_t e : c->ev_dl_part) cudaEventDestroy(e);
  }
  if (c->dl_stage) cudaFree(c->dl_stage);
  if (c->s26_cnt) { cudaFree(c->s26_cnt); cudaFreeHost(c->s26_hcnt); }
  if (c->sg.exec) cudaGraphExecDestroy(c->sg.exec);
  if (c->sg.graph) cudaGraphDestroy(c->sg.graph);
  if (c->sg.d_loop) cudaFree(c->sg.d_loop);
  if (c->sg.h_loop) cudaFreeHost(c->sg.h_loop);
  comm_free(c);
  pair_eam_free(c);
  topo_free(c);
  peratom_free(c);
  langevin_free(c);
  void *ptrs[] = {c->xq, c->xq_tmp, c->cell_xq, c->v, c->v_tmp, c->f, c->xhold, c->tag, c->tag_tmp, c->image,
                  c->image_tmp, c->mass, c->cutneighsq, c->atom_bin, c->cell_count, c->cell_start, c->cell_fill,
                  c->cell_atoms, c->perm, c->scan_block, c->row_off, c->row_sx, c->row_flag, c->neigh, c->numneigh,
                  c->ghost_src, c->ghost_pbc, c->ghost_rank, c->sendbuf, c->recvbuf, c->ljtab, c->rho, c->fp, c->partials,
                  c->d_result, c->d_flags, c->stage, c->tile_ghost, c->tile_pos, c->tiles_int, c->tiles_bnd};
  for (void *p : ptrs) if (p) cudaFree(p);
  if (c->h_result) cudaFreeHost(c->h_result);
  if (c->h_flags) cudaFreeHost(c->h_flags);
  for (auto e : c->ev_pool) cudaEventDestroy(e);
  cudaEvent_t evs[] = {c->ev_x, c->ev_halo, c->ev_fp, c->ev_fphalo, c->ev_bnd, c->ev_rev, c->ev_rb, c->ev_rr, c->ev_clr};
  for (auto e : evs) if (e) cudaEventDestroy(e);
  if (c->comm_stream) cudaStreamDestroy(c->comm_stream);
  if (c->bnd_stream) cudaStreamDestroy(c->bnd_stream);
  cudaStreamDestroy(c->stream);
  delete c;
  return B200MD_OK;
}

extern "C" int b200md_set_units(b200md_ctx *c, double boltz, double mvv2e, double ftm2v, double nktv2p) {
  NEED(c, "b200md: null context");
  c->boltz = boltz; c->mvv2e = mvv2e; c->ftm2v = ftm2v; c->nktv2p = nktv2p;
  return B200MD_OK;
}

extern "C" int b200md_set_box(b200md_ctx *c, const double boxlo[3], const double boxhi[3], const int periodic[3],
                              const int procgrid[3]) {
  NEED(c, "b200md: null context");
  NEED(procgrid[0] * procgrid[1] * procgrid[2] == c->nranks, "Bad grid of processors");
  for (int d = 0; d < 3; d++) {
    NEED(boxhi[d] > boxlo[d], "Box bounds are invalid or missing");
    c->boxlo[d] = boxlo[d]; c->boxhi[d] = boxhi[d]; c->prd[d] = boxhi[d] - boxlo[d];
    c->periodic[d] = periodic[d]; c->procgrid[d] = procgrid[d];
  }
  TRY(b200md_rank2loc(c->rank, c->procgrid, c->myloc));
  TRY(b200md_local_box(c->boxlo, c->boxhi, c->procgrid, c->myloc, c->sublo, c->subhi));
  for (int d = 0; d < 3; d++) {
    int loc[3] = {c->myloc[0], c->myloc[1], c->myloc[2]};
    int left = c->myloc[d] - 1; if (left < 0) left = c->procgrid[d] - 1;
    int right = c->myloc[d] + 1; if (right == c->procgrid[d]) right = 0;
    loc[d] = left; c->procneigh[d][0] = (loc[0] * c->procgrid[1] + loc[1]) * c->procgrid[2] + loc[2];
    loc[d] = right; c->procneigh[d][1] = (loc[0] * c->procgrid[1] + loc[1]) * c->procgrid[2] + loc[2];
  }
  c->box_set = true;
  return B200MD_OK;
}

extern "C" int b200md_get_subbox(b200md_ctx *c, double sublo[3], double subhi[3]) {
  NEED(c && c->box_set, "b200md: box not set");
  for (int d = 0; d < 3; d++) { sublo[d] = c->sublo[d]; subhi[d] = c->subhi[d]; }
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// atoms
// ---------------------------------------------------------------------------------------
__global__ void k_upload_atoms(int n, int nmax, const double *__restrict__ x, const double *__restrict__ v,
                               const int *__restrict__ type, double4 *__restrict__ xq, double *__restrict__ vo,
                               double *__restrict__ f) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  xq[i] = make_double4(x[3 * i], x[3 * i + 1], x[3 * i + 2], __longlong_as_double((long long)type[i]));
  const size_t s = (size_t)nmax;
  vo[i] = v ? v[3 * i] : 0.0; vo[i + s] = v ? v[3 * i + 1] : 0.0; vo[i + 2 * s] = v ? v[3 * i + 2] : 0.0;
  f[i] = 0; f[i + s] = 0; f[i + 2 * s] = 0;
}
__global__ void k_download_atoms(int n, int first, int nmax, const double4 *__restrict__ xq, const double *__restrict__ v,
                                 const double *__restrict__ f, double *__restrict__ x, double *__restrict__ vo,
                                 double *__restrict__ fo, int *__restrict__ type) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int i = first + k;
  const double4 q = xq[i];
  const size_t s = (size_t)nmax;
  if (x) { x[3 * k] = q.x; x[3 * k + 1] = q.y; x[3 * k + 2] = q.z; }
  if (type) type[k] = (int)(__double_as_longlong(q.w) & 0xffffffffll);
  if (vo) { vo[3 * k] = v[i]; vo[3 * k + 1] = v[i + s]; vo[3 * k + 2] = v[i + 2 * s]; }
  if (fo) { fo[3 * k] = f[i]; fo[3 * k + 1] = f[i + s]; fo[3 * k + 2] = f[i + 2 * s]; }
}

extern "C" int b200md_set_atoms(b200md_ctx *c, int nlocal, const double *x, const double *v, const int *type,
                                const int *tag, const int *image, int ntypes, const double *mass) {
  NEED(c && c->box_set, "b200md: box not set");
  NEED(nlocal >= 0 && ntypes >= 1 && x && type && tag && mass, "b200md: bad set_atoms arguments");
  CUDA_TRY(cudaSetDevice(c->device));
  c->nlocal = 0; c->nghost = 0;
  TRY(neigh_grow_atoms(c, nlocal + nlocal / 2 + 4096));
  c->ntypes = ntypes;
  c->h_mass.assign(mass, mass + ntypes + 1);
  for (int t = 1; t <= ntypes; t++) NEED(mass[t] > 0.0, "Not all per-type masses are set");
  if (c->mass) CUDA_TRY(cudaFree(c->mass));
  CUDA_TRY(cudaMalloc(&c->mass, sizeof(double) * (ntypes + 1)));
  CUDA_TRY(cudaMemcpy(c->mass, mass, sizeof(double) * (ntypes + 1), cudaMemcpyHostToDevice));
  if (nlocal) {
    double *dx = nullptr, *dv = nullptr; int *dt = nullptr;
    CUDA_TRY(cudaMalloc(&dx, sizeof(double) * 3 * nlocal));
    CUDA_TRY(cudaMalloc(&dt, sizeof(int) * nlocal));
    CUDA_TRY(cudaMemcpy(dx, x, sizeof(double) * 3 * nlocal, cudaMemcpyHostToDevice));
    CUDA_TRY(cudaMemcpy(dt, type, sizeof(int) * nlocal, cudaMemcpyHostToDevice));
    if (v) {
      CUDA_TRY(cudaMalloc(&dv, sizeof(double) * 3 * nlocal));
      CUDA_TRY(cudaMemcpy(dv, v, sizeof(double) * 3 * nlocal, cudaMemcpyHostToDevice));
    }
    k_upload_atoms<<<(nlocal + 255) / 256, 256, 0, c->stream>>>(nlocal, c->nmax, dx, dv, dt, c->xq, c->v, c->f);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(c->tag, tag, sizeof(int) * nlocal, cudaMemcpyHostToDevice, c->stream));
    if (image) CUDA_TRY(cudaMemcpyAsync(c->image, image, sizeof(int) * nlocal, cudaMemcpyHostToDevice, c->stream));
    else {
      std::vector<int> img(nlocal, (512 << 20) | (512 << 10) | 512);
      CUDA_TRY(cudaMemcpyAsync(c->image, img.data(), sizeof(int) * nlocal, cudaMemcpyHostToDevice, c->stream));
      CUDA_TRY(cudaStreamSynchronize(c->stream));
    }
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    cudaFree(dx); cudaFree(dt); if (dv) cudaFree(dv);
  }
  c->nlocal = nlocal;
  double nat = nlocal;
  TRY(comm_allreduce_sum(c, &nat, 1));
  c->natoms_global = nat;
  c->ago = 0; c->step = 0; c->ev_valid = false;
  c->host_order = true;
  if (c->lang.on) TRY(langevin_rank_from_host_order(c));   // the stream follows the new host order
  return B200MD_OK;
}

extern "C" int b200md_get_nlocal(b200md_ctx *c, int *nlocal, int *nghost) {
  NEED(c, "b200md: null context");
  if (nlocal) *nlocal = c->nlocal;
  if (nghost) *nghost = c->nghost;
  return B200MD_OK;
}

static int stage(b200md_ctx *c, size_t bytes) {
  if (bytes <= c->stage_cap) return B200MD_OK;
  if (c->stage) CUDA_TRY(cudaFree(c->stage));
  c->stage_cap = bytes + bytes / 4 + (1 << 20);
  CUDA_TRY(cudaMalloc(&c->stage, c->stage_cap));
  return B200MD_OK;
}

static int download(b200md_ctx *c, int first, int n, double *x, double *v, double *f, int *type, int *tag, int *image) {
  if (n == 0) return B200MD_OK;
  CUDA_TRY(cudaSetDevice(c->device));
  const size_t n3 = sizeof(double) * 3 * (size_t)n;
  TRY(stage(c, 3 * n3 + sizeof(int) * (size_t)n));
  double *dx = x ? (double *)c->stage : nullptr;
  double *dv = v ? (double *)(c->stage + n3) : nullptr;
  double *df = f ? (double *)(c->stage + 2 * n3) : nullptr;
  int *dt = type ? (int *)(c->stage + 3 * n3) : nullptr;
  k_download_atoms<<<(n + 255) / 256, 256, 0, c->stream>>>(n, first, c->nmax, c->xq, c->v, c->f, dx, dv, df, dt);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  if (x) CUDA_TRY(cudaMemcpyAsync(x, dx, n3, cudaMemcpyDeviceToHost, c->stream));
  if (v) CUDA_TRY(cudaMemcpyAsync(v, dv, n3, cudaMemcpyDeviceToHost, c->stream));
  if (f) CUDA_TRY(cudaMemcpyAsync(f, df, n3, cudaMemcpyDeviceToHost, c->stream));
  if (type) CUDA_TRY(cudaMemcpyAsync(type, dt, sizeof(int) * n, cudaMemcpyDeviceToHost, c->stream));
  if (tag) CUDA_TRY(cudaMemcpyAsync(tag, c->tag + first, sizeof(int) * n, cudaMemcpyDeviceToHost, c->stream));
  if (image) CUDA_TRY(cudaMemcpyAsync(image, c->image + first, sizeof(int) * n, cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  return B200MD_OK;
}

__global__ void k_put_xv(int n, int nmax, const double *__restrict__ x, const double *__restrict__ v,
                         double4 *__restrict__ xq, double *__restrict__ vo) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (x) { double4 q = xq[i]; q.x = x[3 * i]; q.y = x[3 * i + 1]; q.z = x[3 * i + 2]; xq[i] = q; }
  if (v) { const size_t s = (size_t)nmax; vo[i] = v[3 * i]; vo[i + s] = v[3 * i + 1]; vo[i + 2 * s] = v[3 * i + 2]; }
}

extern "C" int b200md_put_atoms(b200md_ctx *c, const double *x, const double *v) {
  NEED(c, "b200md: null context");
  const int n = c->nlocal;
  if (n == 0 || (!x && !v)) return B200MD_OK;
  CUDA_TRY(cudaSetDevice(c->device));
  const size_t n3 = sizeof(double) * 3 * (size_t)n;
  TRY(stage(c, 2 * n3));
  double *dx = x ? (double *)c->stage : nullptr, *dv = v ? (double *)(c->stage + n3) : nullptr;
  if (x) CUDA_TRY(cudaMemcpyAsync(dx, x, n3, cudaMemcpyHostToDevice, c->stream));
  if (v) CUDA_TRY(cudaMemcpyAsync(dv, v, n3, cudaMemcpyHostToDevice, c->stream));
  k_put_xv<<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->nmax, dx, dv, c->xq, c->v);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

extern "C" int b200md_host_alloc(void **ptr, size_t bytes) {
  CUDA_TRY(cudaMallocHost(ptr, bytes));
  return B200MD_OK;
}
extern "C" int b200md_host_free(void *ptr) {
  CUDA_TRY(cudaFreeHost(ptr));
  return B200MD_OK;
}

// Asynchronous download for a host that reads results every step (bench.py e2e): x and f of the owned atoms are packed
// on the context's stream into a device staging area of their own, then copied on a COPY stream, x first.  The caller
// waits for x (b200md_wait_download(ctx, 0)), may upload the next step's input at once, and the f copy (D2H) runs
// while that upload (H2D, the other PCIe direction) and the next step's kernels are in flight.
__global__ void k_pack_xf(int n, int nmax, const double4 *__restrict__ xq, const double *__restrict__ f,
                          double *__restrict__ x, double *__restrict__ fo) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const size_t s = (size_t)nmax;
  if (x) { const double4 q = xq[i]; x[3 * i] = q.x; x[3 * i + 1] = q.y; x[3 * i + 2] = q.z; }
  if (fo) { fo[3 * i] = f[i]; fo[3 * i + 1] = f[i + s]; fo[3 * i + 2] = f[i + 2 * s]; }
}
extern "C" int b200md_get_atoms_async(b200md_ctx *c, double *x, double *f) {
  NEED(c, "b200md: null context");
  const int n = c->nlocal;
  if (n == 0 || (!x && !f)) return B200MD_OK;
  CUDA_TRY(cudaSetDevice(c->device));
  if (!c->copy_stream) {
    CUDA_TRY(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    CUDA_TRY(cudaEventCreateWithFlags(&c->ev_packed, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&c->ev_dl_x, cudaEventDisableTiming));
    CUDA_TRY(cudaEventCreateWithFlags(&c->ev_dl_all, cudaEventDisableTiming));
  }
  const size_t n3 = sizeof(double) * 3 * (size_t)n;
  if (2 * n3 > c->dl_cap) {
    CUDA_TRY(cudaStreamSynchronize(c->copy_stream));
    if (c->dl_stage) CUDA_TRY(cudaFree(c->dl_stage));
    c->dl_cap = 2 * n3 + n3 / 4 + (1 << 20);
    CUDA_TRY(cudaMalloc(&c->dl_stage, c->dl_cap));
  }
  double *dx = x ? (double *)c->dl_stage : nullptr, *df = f ? (double *)(c->dl_stage + n3) : nullptr;
  if (c->dl_pending) CUDA_TRY(cudaStreamWaitEvent(c->stream, c->ev_dl_all, 0));   // the staging area is free again
  k_pack_xf<<<(n + 255) / 256, 256, 0, c->stream>>>(n, c->nmax, c->xq, c->f, dx, df);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  CUDA_TRY(cudaEventRecord(c->ev_packed, c->stream));
  CUDA_TRY(cudaStreamWaitEvent(c->copy_stream, c->ev_packed, 0));
  // x in B200MD_DL_PARTS pieces of whole atoms, an event behind each: b200md_wait_download_part lets the host turn a
  // piece around (b200md_put_atoms_range) while the next one is still on the link
  c->dl_n = n;
  c->dl_parts = x ? std::min(B200MD_DL_PARTS, std::max(1, n / 65536)) : 0;
  while ((int)c->ev_dl_part.size() < c->dl_parts) {
    cudaEvent_t e;
    CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    c->ev_dl_part.push_back(e);
  }
  for (int p = 0; p < c->dl_parts; p++) {
    const size_t a0 = (size_t)n * p / c->dl_parts, a1 = (size_t)n * (p + 1) / c->dl_parts;
    CUDA_TRY(cudaMemcpyAsync(x + 3 * a0, dx + 3 * a0, sizeof(double) * 3 * (a1 - a0), cudaMemcpyDeviceToHost, c->copy_stream));
    CUDA_TRY(cudaEventRecord(c->ev_dl_part[p], c->copy_stream));
  }
  CUDA_TRY(cudaEventRecord(c->ev_dl_x, c->copy_stream));
  if (f) CUDA_TRY(cudaMemcpyAsync(f, df, n3, cudaMemcpyDeviceToHost, c->copy_stream));
  CUDA_TRY(cudaEventRecord(c->ev_dl_all, c->copy_stream));
  c->dl_pending = true;
  return B200MD_OK;
}
extern "C" int b200md_wait_download(b200md_ctx *c, int all) {
  NEED(c, "b200md: null context");
  if (!c->dl_pending) return B200MD_OK;
  CUDA_TRY(cudaEventSynchronize(all ? c->ev_dl_all : c->ev_dl_x));
  if (all) c->dl_pending = false;
  return B200MD_OK;
}

extern "C" int b200md_wait_download_part(b200md_ctx *c, int part, int *first, int *count) {
  NEED(c && first && count, "b200md: null argument");
  *first = 0; *count = 0;
  if (!c->dl_pending || part < 0 || part >= c->dl_parts) return B200MD_OK;
  CUDA_TRY(cudaEventSynchronize(c->ev_dl_part[part]));
  const size_t a0 = (size_t)c->dl_n * part / c->dl_parts, a1 = (size_t)c->dl_n * (part + 1) / c->dl_parts;
  *first = (int)a0; *count = (int)(a1 - a0);
  return B200MD_OK;
}

// positions of the owned atoms [first, first+count) from the host array x (row 0 = atom 0)
__global__ void k_put_x_range(int first, int count, const double *__restrict__ x, double4 *__restrict__ xq) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  i += first;
  double4 q = xq[i]; q.x = x[3 * i]; q.y = x[3 * i + 1]; q.z = x[3 * i + 2]; xq[i] = q;
}
extern "C" int b200md_put_atoms_range(b200md_ctx *c, const double *x, int first, int count) {
  NEED(c && x, "b200md: null argument");
  NEED(first >= 0 && count >= 0 && first + count <= c->nlocal, "b200md_put_atoms_range: range outside the owned atoms");
  if (count == 0) return B200MD_OK;
  CUDA_TRY(cudaSetDevice(c->device));
  TRY(stage(c, 2 * sizeof(double) * 3 * (size_t)c->nlocal));     // same staging area and size as b200md_put_atoms
  double *dx = (double *)c->stage;
  CUDA_TRY(cudaMemcpyAsync(dx + 3 * (size_t)first, x + 3 * (size_t)first, sizeof(double) * 3 * (size_t)count,
                           cudaMemcpyHostToDevice, c->stream));
  k_put_x_range<<<(count + 255) / 256, 256, 0, c->stream>>>(first, count, dx, c->xq);
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

extern "C" int b200md_get_atoms(b200md_ctx *c, double *x, double *v, double *f, int *type, int *tag, int *image) {
  NEED(c, "b200md: null context");
  return download(c, 0, c->nlocal, x, v, f, type, tag, image);
}
extern "C" int b200md_get_ghosts(b200md_ctx *c, double *x, int *type, int *tag) {
  NEED(c, "b200md: null context");
  return download(c, c->nlocal, c->nghost, x, nullptr, nullptr, type, tag, nullptr);
}

extern "C" int b200md_velocity_finalize(b200md_ctx *c, double t_desired, int zero_momentum) {
  NEED(c, "b200md: null context");
  CUDA_TRY(cudaSetDevice(c->device));
  double s[5];
  double sx = 0, sy = 0, sz = 0;
  if (zero_momentum) {
    // Velocity::zero_momentum: subtract the centre-of-mass velocity (src/velocity.cpp, group->vcm)
    TRY(thermo_ke(c, s));
    TRY(comm_allreduce_sum(c, s, 5));
    if (s[4] > 0) { sx = s[1] / s[4]; sy = s[2] / s[4]; sz = s[3] / s[4]; }
    TRY(vel_adjust(c, sx, sy, sz, 1.0));
  }
  // Velocity::rescale: factor = sqrt(t_new/t_old) with ComputeTemp (dof = 3N-3)
  TRY(thermo_ke(c, s));
  TRY(comm_allreduce_sum(c, s, 5));
  const double dof = 3.0 * c->natoms_global - 3.0;
  const double tfactor = dof > 0 ? c->mvv2e / (dof * c->boltz) : 0.0;
  const double t_old = s[0] * tfactor;
  NEED(t_old != 0.0, "Attempting to rescale a 0.0 temperature");
  TRY(vel_adjust(c, 0, 0, 0, sqrt(t_desired / t_old)));
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// neighbor
// ---------------------------------------------------------------------------------------
extern "C" int b200md_neigh_config(b200md_ctx *c, int ntypes, const double *cutneighsq, double cutneighmax,
                                   double skin, int every, int delay, int check, int list_kind) {
  NEED(c && c->box_set, "b200md: box not set");
  NEED(ntypes >= 1 && cutneighsq && cutneighmax > 0, "b200md: bad neigh_config arguments");
  NEED(every >= 1 && delay >= 0, "Illegal neigh_modify command");
  NEED(list_kind == B200MD_LIST_FULL || list_kind == B200MD_LIST_HALF_NEWTON, "b200md: unknown list kind");
  CUDA_TRY(cudaSetDevice(c->device));
  const size_t nt = (size_t)(ntypes + 1) * (ntypes + 1);
  c->h_cutneighsq.assign(cutneighsq, cutneighsq + nt);
  if (c->cutneighsq) CUDA_TRY(cudaFree(c->cutneighsq));
  CUDA_TRY(cudaMalloc(&c->cutneighsq, sizeof(double) * nt));
  CUDA_TRY(cudaMemcpy(c->cutneighsq, cutneighsq, sizeof(double) * nt, cudaMemcpyHostToDevice));
  c->cutneighmax = cutneighmax;
  c->cutghost = cutneighmax;  // Comm::get_comm_cutoff: max(cutneighmax, user cutoff=0) (src/comm.cpp:681-730)
  c->skin = skin;
  c->triggersq = 0.25 * skin * skin;  // src/neighbor.cpp:348
  c->every = every; c->delay = delay; c->check = check;
  c->list_kind = list_kind;
  c->pitch = 0;
  TRY(comm_setup(c));
  TRY(neigh_setup_bins(c));
  c->neigh_configured = true;
  return B200MD_OK;
}

static bool decide_will_check(const b200md_ctx *c) {
  const int ago = c->ago + 1;
  return ago >= c->delay && ago % c->every == 0 && c->check;
}

extern "C" int b200md_neigh_decide(b200md_ctx *c, int *rebuild) {
  NEED(c && c->neigh_configured, "b200md: neighbor not configured");
  c->ago++;
  *rebuild = 0;
  if (c->ago >= c->delay && c->ago % c->every == 0) {
    if (!c->check) { *rebuild = 1; return B200MD_OK; }
    // check_distance: the flag was raised by the fused initial-integrate kernel
    CUDA_TRY(cudaMemcpyAsync(c->h_flags, c->d_flags, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    int flag = c->h_flags[0];
    TRY(comm_allreduce_max_int(c, &flag));
    if (flag && c->ago == (c->every > c->delay ? c->every : c->delay)) c->ndanger++;
    *rebuild = flag;
  }
  return B200MD_OK;
}

extern "C" int b200md_neigh_build(b200md_ctx *c) {
  NEED(c && c->neigh_configured, "b200md: neighbor not configured");
  CUDA_TRY(cudaSetDevice(c->device));
  TRY(neigh_choose_layout(c));
  TRY(neigh_pbc(c));
  TRY(comm_exchange(c));
  if (c->lang.on && c->lang.resort_pending) {   // Verlet::run's Atom::sort (verlet.cpp:285): the draws follow the new order
    c->lang.resort_pending = false;
    TRY(langevin_resort(c));
  }
  c->host_order = false;
  TRY(neigh_sort_owned(c));
  TRY(comm_borders(c));
  TRY(neigh_bin_all(c));
  TRY(neigh_build_list(c));
  TRY(topo_build_map(c));
  TRY(neigh_save_xhold(c));
  CUDA_TRY(cudaMemsetAsync(c->d_flags, 0, sizeof(int), c->stream));
  c->ago = 0;
  c->nbuilds++;
  return B200MD_OK;
}

extern "C" int b200md_forward_comm(b200md_ctx *c) {
  NEED(c && c->neigh_configured, "b200md: neighbor not configured");
  return comm_forward(c);
}
extern "C" int b200md_reverse_comm(b200md_ctx *c) {
  NEED(c && c->neigh_configured, "b200md: neighbor not configured");
  return comm_reverse_force(c);
}

__global__ void k_export_list(int nlocal, int tpa, int pitch, const int *__restrict__ neigh, const int *__restrict__ numneigh,
                              const long long *__restrict__ offs, const int *__restrict__ tag, int by_tag, int *__restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const int A = 32 / tpa;
  const int tile = i / A, sub = i % A;
  const int *nb = neigh + (size_t)tile * pitch * A + sub * tpa;
  const int n = numneigh[i];
  int *o = out + offs[i];
  for (int k = 0; k < n; k++) {
    const int j = nb[(k / tpa) * 32 + (k % tpa)];
    o[k] = by_tag ? tag[j] : j;
  }
}

// cluster-pair list -> the exact per-atom lists: atom k of a cluster owns the entries whose mask has bit k
__global__ void k_cluster_count(int nlocal, int tpa, int pitch, const int *__restrict__ neigh, const int *__restrict__ numneigh,
                                int *__restrict__ count) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const int c = i / CL, k = i % CL, C = 32 / tpa;
  const int *nb = neigh + (size_t)(c / C) * pitch * C + (c % C) * tpa;
  const int n = numneigh[c];
  int cnt = 0;
  for (int s = 0; s < n; s++) cnt += ((unsigned)nb[(s / tpa) * 32 + (s % tpa)] >> (CL_SHIFT + k)) & 1u;
  count[i] = cnt;
}
__global__ void k_export_cluster(int nlocal, int tpa, int pitch, const int *__restrict__ neigh, const int *__restrict__ numneigh,
                                 const long long *__restrict__ offs, const int *__restrict__ tag, int by_tag, int *__restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const int c = i / CL, k = i % CL, C = 32 / tpa;
  const int *nb = neigh + (size_t)(c / C) * pitch * C + (c % C) * tpa;
  const int n = numneigh[c];
  int *o = out + offs[i];
  int w = 0;
  for (int s = 0; s < n; s++) {
    const unsigned e = (unsigned)nb[(s / tpa) * 32 + (s % tpa)];
    if ((e >> (CL_SHIFT + k)) & 1u) { const int j = (int)(e & CL_JMASK); o[w++] = by_tag ? tag[j] : j; }
  }
}

extern "C" int b200md_neigh_export(b200md_ctx *c, int by_tag, int *numneigh, int *neigh_flat, long long *total) {
  NEED(c && c->neigh && c->nbuilds > 0, "b200md: no neighbor list built");
  CUDA_TRY(cudaSetDevice(c->device));
  const int n = c->nlocal;
  std::vector<int> nn(n);
  int *d_count = nullptr;
  if (c->cluster && n) {
    CUDA_TRY(cudaMalloc(&d_count, sizeof(int) * n));
    k_cluster_count<<<(n + 127) / 128, 128, 0, c->stream>>>(n, c->tpa, c->pitch, c->neigh, c->numneigh, d_count);
    CUDA_TRY(cudaGetLastError());
  }
  CUDA_TRY(cudaMemcpyAsync(nn.data(), c->cluster ? d_count : c->numneigh, sizeof(int) * n, cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  if (d_count) cudaFree(d_count);
  std::vector<long long> offs(n + 1);
  long long t = 0;
  for (int i = 0; i < n; i++) { offs[i] = t; t += nn[i]; }
  offs[n] = t;
  if (numneigh) memcpy(numneigh, nn.data(), sizeof(int) * n);
  if (total) *total = t;
  if (neigh_flat && t > 0) {
    long long *d_offs = nullptr; int *d_out = nullptr;
    CUDA_TRY(cudaMalloc(&d_offs, sizeof(long long) * (n + 1)));
    CUDA_TRY(cudaMalloc(&d_out, sizeof(int) * t));
    CUDA_TRY(cudaMemcpy(d_offs, offs.data(), sizeof(long long) * (n + 1), cudaMemcpyHostToDevice));
    if (c->cluster)
      k_export_cluster<<<(n + 127) / 128, 128, 0, c->stream>>>(n, c->tpa, c->pitch, c->neigh, c->numneigh, d_offs, c->tag, by_tag, d_out);
    else
      k_export_list<<<(n + 127) / 128, 128, 0, c->stream>>>(n, c->tpa, c->pitch, c->neigh, c->numneigh, d_offs, c->tag, by_tag, d_out);
    CUDA_TRY(cudaGetLastError());
    CUDA_TRY(cudaMemcpyAsync(neigh_flat, d_out, sizeof(int) * t, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    cudaFree(d_offs); cudaFree(d_out);
  }
  return B200MD_OK;
}

extern "C" int b200md_neigh_stats(b200md_ctx *c, long long *nbuilds, long long *ndanger, long long *total_neighbors,
                                  int *maxneigh_pitch) {
  NEED(c, "b200md: null context");
  if (nbuilds) *nbuilds = c->nbuilds;
  if (ndanger) *ndanger = c->ndanger;
  if (maxneigh_pitch) *maxneigh_pitch = c->pitch;
  if (total_neighbors) {
    *total_neighbors = 0;
    if (c->nbuilds > 0) TRY(b200md_neigh_export(c, 0, nullptr, nullptr, total_neighbors));
  }
  return B200MD_OK;
}

// storage layout of the current list: per-atom rows or cluster rows (4 atoms share a row whose entries carry a 4-bit
// pair mask), the number of stored entries and the lanes that walk a row
extern "C" int b200md_neigh_layout(b200md_ctx *c, int *cluster, long long *entries, int *lanes) {
  NEED(c && c->neigh && c->nbuilds > 0, "b200md: no neighbor list built");
  CUDA_TRY(cudaSetDevice(c->device));
  if (cluster) *cluster = c->cluster ? 1 : 0;
  if (lanes) *lanes = c->tpa;
  if (entries) {
    const int n = list_units(c);
    std::vector<int> nn(n);
    if (n) CUDA_TRY(cudaMemcpyAsync(nn.data(), c->numneigh, sizeof(int) * n, cudaMemcpyDeviceToHost, c->stream));
    CUDA_TRY(cudaStreamSynchronize(c->stream));
    long long t = 0;
    for (int v : nn) t += v;
    *entries = t;
  }
  return B200MD_OK;
}

extern "C" int b200md_set_lanes_per_atom(b200md_ctx *c, int lanes) {
  NEED(c, "b200md: null context");
  NEED(lanes == 1 || lanes == 2 || lanes == 4 || lanes == 8 || lanes == 16 || lanes == 32,
       "b200md: lanes per atom must be 1, 2, 4, 8, 16 or 32");
  c->tpa = lanes;
  c->tpa_user = true;
  return B200MD_OK;
}
extern "C" int b200md_get_stream(b200md_ctx *c, void **stream) {
  NEED(c && stream, "b200md: null argument");
  *stream = (void *)c->stream;
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// runtime options: what the reference's accelerator packages take through the `package` command
// (src/input.cpp Input::package -> FixGPU / KokkosLMP::accelerator, src/KOKKOS/kokkos.cpp:300-560)
// ---------------------------------------------------------------------------------------
static bool opt_bool(const char *v, bool *out) {
  if (!strcmp(v, "1") || !strcmp(v, "on") || !strcmp(v, "yes")) { *out = true; return true; }
  if (!strcmp(v, "0") || !strcmp(v, "off") || !strcmp(v, "no")) { *out = false; return true; }
  return false;
}
static bool lanes_ok(int l) { return l == 1 || l == 2 || l == 4 || l == 8 || l == 16 || l == 32; }

extern "C" int b200md_set_option(b200md_ctx *c, const char *key, const char *value) {
  NEED(c && key && value, "b200md: null argument");
  const std::string k(key);
  bool b = false;
  if (k == "halo") {
    if (!strcmp(value, "direct")) c->halo_mode = 2;
    else if (!strcmp(value, "owner")) c->halo_mode = 1;
    else if (!strcmp(value, "staged")) c->halo_mode = 0;
    else { b200md_set_error("Illegal package b200 command: halo must be direct, owner or staged"); return B200MD_ERR; }
    if (c->neigh_configured) TRY(comm_setup(c));
  } else if (k == "overlap") {
    NEED(opt_bool(value, &b), "Illegal package b200 command: overlap must be on or off");
    c->overlap = b;
  } else if (k == "split") {
    const double s = atof(value);
    NEED(s >= 0.0 && s <= 1.0, "Illegal package b200 command: split must be in [0,1]");
    c->split = s;
  } else if (k == "fuse_nve") {
    NEED(opt_bool(value, &b), "Illegal package b200 command: fuse_nve must be on or off");
    c->fuse_nve = b;
  } else if (k == "eam_smem") {
    if (!strcmp(value, "auto")) c->eam_smem = -1;
    else {
      NEED(opt_bool(value, &b), "Illegal package b200 command: eam_smem must be on, off or auto");
      c->eam_smem = b ? 1 : 0;
    }
  } else if (k == "self_border") {
    if (!strcmp(value, "direct")) c->self26 = true;
    else if (!strcmp(value, "staged")) c->self26 = false;
    else { b200md_set_error("Illegal package b200 command: self_border must be direct or staged"); return B200MD_ERR; }
  } else if (k == "cell_order") {
    if (!strcmp(value, "x")) c->cell_by_x = true;
    else if (!strcmp(value, "index")) c->cell_by_x = false;
    else { b200md_set_error("Illegal package b200 command: cell_order must be x or index"); return B200MD_ERR; }
  } else if (k == "bin_scale") {
    if (!strcmp(value, "auto")) c->bin_scale_user = 0;
    else {
      const double v = atof(value);
      NEED(v >= 0.25 && v <= 4.0, "Illegal package b200 command: bin_scale must be auto or in [0.25,4]");
      c->bin_scale_user = v;
    }
  } else if (k == "eam_mask") {
    NEED(opt_bool(value, &b), "Illegal package b200 command: eam_mask must be on or off");
    c->eam_mask = b;
  } else if (k == "cluster") {
    if (!strcmp(value, "auto")) c->use_cluster = -1;
    else {
      NEED(opt_bool(value, &b), "Illegal package b200 command: cluster must be on, off or auto");
      c->use_cluster = b ? 1 : 0;
    }
  } else if (k == "graph") {
    if (!strcmp(value, "auto")) c->sg.mode = -1;
    else {
      NEED(opt_bool(value, &b), "Illegal package b200 command: graph must be on, off or auto");
      c->sg.mode = b ? 1 : 0;
    }
  } else if (k == "trace") {
    NEED(opt_bool(value, &b), "Illegal package b200 command: trace must be on or off");
    c->trace = b;
  } else if (k == "phase_timing") {
    NEED(opt_bool(value, &b), "Illegal package b200 command: phase_timing must be on or off");
    c->phase_timing = b;
  } else if (k == "build_lanes") {
    const int l = atoi(value);
    NEED(lanes_ok(l), "Illegal package b200 command: build_lanes must be 1, 2, 4, 8, 16 or 32");
    c->tpa_build = l; c->tpa_build_user = true;
  } else if (k == "lanes") {
    return b200md_set_lanes_per_atom(c, atoi(value));
  } else {
    b200md_set_error("Illegal package b200 command: unknown keyword " + k);
    return B200MD_ERR;
  }
  return B200MD_OK;
}

extern "C" int b200md_get_option(b200md_ctx *c, const char *key, char *value, int len) {
  NEED(c && key && value && len > 0, "b200md: null argument");
  const std::string k(key);
  std::string v;
  if (k == "halo") v = c->nranks == 1 ? "none" : (c->h26.on || (!c->neigh_configured && c->halo_mode == 2) ? "direct" : (c->direct.on || c->halo_mode == 1 ? "owner" : "staged"));
  else if (k == "overlap") v = c->overlap ? "on" : "off";
  else if (k == "split") v = std::to_string(c->split);
  else if (k == "fuse_nve") v = c->fuse_nve ? "on" : "off";
  else if (k == "eam_smem") v = c->eam_smem < 0 ? "auto" : (c->eam_smem ? "on" : "off");
  else if (k == "eam_mask") v = c->eam_mask ? "on" : "off";
  else if (k == "bin_scale") v = c->bin_scale_user > 0 ? std::to_string(c->bin_scale_user) : "auto";
  else if (k == "cell_order") v = c->cell_by_x ? "x" : "index";
  else if (k == "self_border") v = c->self26 ? "direct" : "staged";
  else if (k == "graph") v = c->sg.mode < 0 ? "auto" : (c->sg.mode ? "on" : "off");
  else if (k == "cluster") v = c->use_cluster < 0 ? "auto" : (c->use_cluster ? "on" : "off");
  else if (k == "phase_timing") v = c->phase_timing ? "on" : "off";
  else if (k == "build_lanes") v = std::to_string(c->tpa_build);
  else if (k == "lanes") v = std::to_string(c->tpa);
  else { b200md_set_error("Illegal package b200 command: unknown keyword " + k); return B200MD_ERR; }
  snprintf(value, len, "%s", v.c_str());
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// pair
// ---------------------------------------------------------------------------------------
extern "C" int b200md_pair_lj_cut_config(b200md_ctx *c, int ntypes, const double *cutsq, const double *lj1,
                                         const double *lj2, const double *lj3, const double *lj4, const double *offset) {
  NEED(c, "b200md: null context");
  NEED(ntypes >= 1 && cutsq && lj1 && lj2 && lj3 && lj4 && offset, "All pair coeffs are not set");
  NEED(c->ntypes == 0 || c->ntypes == ntypes, "b200md: ntypes differs from set_atoms");
  CUDA_TRY(cudaSetDevice(c->device));
  const size_t nt = (size_t)(ntypes + 1) * (ntypes + 1);
  c->h_ljtab.assign(nt * 8, 0.0);
  double cmax = 0;
  for (size_t k = 0; k < nt; k++) {
    double *t = &c->h_ljtab[k * 8];
    t[0] = cutsq[k]; t[1] = lj1[k]; t[2] = lj2[k]; t[3] = lj3[k]; t[4] = lj4[k]; t[5] = offset[k];
    if (cutsq[k] > cmax) cmax = cutsq[k];
  }
  c->cutforcemax = sqrt(cmax);
  if (c->ljtab) CUDA_TRY(cudaFree(c->ljtab));
  CUDA_TRY(cudaMalloc(&c->ljtab, sizeof(double) * nt * 8));
  CUDA_TRY(cudaMemcpy(c->ljtab, c->h_ljtab.data(), sizeof(double) * nt * 8, cudaMemcpyHostToDevice));
  c->pair_style = B200MD_PAIR_LJ_CUT;
  return B200MD_OK;
}

extern "C" int b200md_pair_eam_config(b200md_ctx *c, int ntypes, int nrho, int nr, double rdr, double rdrho,
                                      double rhomax, double cutforcesq, int nfrho, const double *frho_spline,
                                      int nrhor, const double *rhor_spline, int nz2r, const double *z2r_spline,
                                      const int *type2frho, const int *type2rhor, const int *type2z2r,
                                      const double *scale) {
  NEED(c, "b200md: null context");
  NEED(ntypes >= 1 && frho_spline && rhor_spline && z2r_spline && type2frho && type2rhor && type2z2r,
       "All pair coeffs are not set");
  CUDA_TRY(cudaSetDevice(c->device));
  TRY(pair_eam_config_impl(c, ntypes, nrho, nr, rdr, rdrho, rhomax, cutforcesq, nfrho, frho_spline, nrhor,
                           rhor_spline, nz2r, z2r_spline, type2frho, type2rhor, type2z2r, scale));
  c->cutforcemax = sqrt(cutforcesq);
  c->pair_style = B200MD_PAIR_EAM;
  return B200MD_OK;
}

extern "C" int b200md_pair_compute(b200md_ctx *c, int eflag, int vflag, double *eng_vdwl, double virial[6]) {
  NEED(c && c->nbuilds > 0, "b200md: pair_compute before the first neighbor build");
  CUDA_TRY(cudaSetDevice(c->device));
  if (c->pair_style == B200MD_PAIR_LJ_CUT) TRY(pair_lj_compute(c, eflag, vflag));
  else if (c->pair_style == B200MD_PAIR_EAM) TRY(pair_eam_compute(c, eflag, vflag));
  else { b200md_set_error("b200md: no pair style configured"); return B200MD_ERR; }
  if (eflag && eng_vdwl) *eng_vdwl = c->eng_vdwl;
  if (vflag && virial) for (int k = 0; k < 6; k++) virial[k] = c->virial[k];
  return B200MD_OK;
}

extern "C" int b200md_pair_tallies(b200md_ctx *c, double *eng_vdwl, double virial[6]) {
  NEED(c, "b200md: null context");
  NEED(c->ev_valid, "Energy was not tallied on needed timestep");
  if (eng_vdwl) *eng_vdwl = c->eng_vdwl;
  if (virial) for (int k = 0; k < 6; k++) virial[k] = c->virial[k];
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// molecular topology, bonds
// ---------------------------------------------------------------------------------------
extern "C" int b200md_set_topology(b200md_ctx *c, long long natoms, const int *bond_offset, const int *bond_partner_tag,
                                   const int *bond_type, const int *special_offset, const int *special_tag,
                                   const int *special_level, const double special_lj[4], const double special_coul[4]) {
  NEED(c && natoms > 0 && bond_offset && special_offset && special_lj && special_coul, "b200md: null argument");
  NEED(natoms < 2147483647LL, "b200md: tags must fit 32 bits");
  CUDA_TRY(cudaSetDevice(c->device));
  return topo_set(c, natoms, bond_offset, bond_partner_tag, bond_type, special_offset, special_tag, special_level,
                  special_lj, special_coul);
}
extern "C" int b200md_bond_fene_config(b200md_ctx *c, int nbondtypes, const double *k, const double *r0,
                                       const double *epsilon, const double *sigma) {
  NEED(c && nbondtypes >= 1 && k && r0 && epsilon && sigma, "All bond coeffs are not set");
  CUDA_TRY(cudaSetDevice(c->device));
  return bond_fene_set(c, nbondtypes, k, r0, epsilon, sigma);
}
extern "C" int b200md_bond_compute(b200md_ctx *c, int eflag, int vflag, double *energy, double virial[6]) {
  NEED(c && c->nbuilds > 0, "b200md: bond_compute before the first neighbor build");
  NEED(c->topo.molecular && c->topo.fene, "b200md: no topology / bond style configured");
  CUDA_TRY(cudaSetDevice(c->device));
  TRY(bond_compute(c, eflag, vflag));
  if (eflag && energy) *energy = c->topo.eng_bond;
  if (vflag && virial) for (int k = 0; k < 6; k++) virial[k] = c->topo.virial_bond[k];
  return B200MD_OK;
}

extern "C" int b200md_bond_tallies(b200md_ctx *c, double *energy, double virial[6]) {
  NEED(c, "b200md: null context");
  NEED(c->ev_valid, "Energy was not tallied on needed timestep");
  if (energy) *energy = c->topo.eng_bond;
  if (virial) for (int k = 0; k < 6; k++) virial[k] = c->topo.virial_bond[k];
  return B200MD_OK;
}

// per-atom energy / virial of the current configuration (peratom.cu)
extern "C" int b200md_peratom(b200md_ctx *c, int eflag_atom, int vflag_atom, int which, double *eatom, double *vatom) {
  NEED(c && c->neigh_configured && c->nbuilds > 0, "b200md: per-atom tallies before setup");
  NEED(which & (B200MD_PERATOM_PAIR | B200MD_PERATOM_BOND), "b200md: nothing selected for the per-atom tallies");
  NEED(!(eflag_atom && !eatom) && !(vflag_atom && !vatom), "b200md: null output for a requested per-atom tally");
  CUDA_TRY(cudaSetDevice(c->device));
  TRY(peratom_compute(c, eflag_atom, vflag_atom, which));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  const size_t n = (size_t)c->nlocal;
  if (eflag_atom && n) CUDA_TRY(cudaMemcpy(eatom, c->peratom, sizeof(double) * n, cudaMemcpyDeviceToHost));
  if (vflag_atom && n) {
    std::vector<double> col(n);
    for (int k = 0; k < 6; k++) {
      CUDA_TRY(cudaMemcpy(col.data(), c->peratom + (size_t)(k + 1) * c->nmax, sizeof(double) * n, cudaMemcpyDeviceToHost));
      for (size_t i = 0; i < n; i++) vatom[i * 6 + k] = col[i];
    }
  }
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// fix langevin (langevin.cu)
// ---------------------------------------------------------------------------------------
extern "C" int b200md_fix_langevin(b200md_ctx *c, double t_start, double t_stop, double t_period, int seed,
                                   const double *ratio, double dt) {
  NEED(c, "b200md: null context");
  CUDA_TRY(cudaSetDevice(c->device));
  return langevin_config(c, t_start, t_stop, t_period, seed, ratio, dt);
}
extern "C" int b200md_fix_langevin_window(b200md_ctx *c, long long ntimestep, long long beginstep, long long endstep) {
  NEED(c && c->lang.on, "b200md: fix langevin is not set");
  NEED(endstep >= beginstep && ntimestep >= beginstep && (ntimestep <= endstep || endstep == beginstep),
       "b200md: bad fix langevin window");
  Langevin &L = c->lang;
  L.window = endstep > beginstep;
  L.begin = beginstep; L.end = endstep; L.now = ntimestep;
  return B200MD_OK;
}
extern "C" int b200md_fix_langevin_sort(b200md_ctx *c, int sortfreq) {
  NEED(c && c->lang.on, "b200md: fix langevin is not set");
  NEED(sortfreq >= 0, "Illegal atom_modify sort frequency");
  c->lang.sortfreq = sortfreq;
  return B200MD_OK;
}
extern "C" int b200md_fix_langevin_off(b200md_ctx *c) {
  NEED(c, "b200md: null context");
  c->lang.on = false;
  return B200MD_OK;
}
extern "C" int b200md_ranmars_device_stream(b200md_ctx *c, int seed, int n, int fills, double *out) {
  NEED(c && n >= 0 && fills >= 0 && (out || !n || !fills), "b200md: bad ranmars_device_stream arguments");
  CUDA_TRY(cudaSetDevice(c->device));
  return langevin_test_stream(c, seed, n, fills, out);
}

// ---------------------------------------------------------------------------------------
// fix nve, thermo
// ---------------------------------------------------------------------------------------
extern "C" int b200md_nve_initial(b200md_ctx *c, double dtv, double dtf) {
  NEED(c && c->neigh_configured, "b200md: neighbor not configured");
  CUDA_TRY(cudaSetDevice(c->device));
  return nve_initial(c, dtv, dtf, decide_will_check(c) && c->nbuilds > 0);
}
extern "C" int b200md_nve_final(b200md_ctx *c, double dtf) {
  NEED(c, "b200md: null context");
  CUDA_TRY(cudaSetDevice(c->device));
  return nve_final(c, dtf);
}

extern "C" int b200md_thermo_compute(b200md_ctx *c, b200md_thermo *out) {
  NEED(c && out, "b200md: null argument");
  NEED(c->ev_valid, "Energy was not tallied on needed timestep");
  CUDA_TRY(cudaSetDevice(c->device));
  double s[5];
  TRY(thermo_ke(c, s));
  const Topology &tp = c->topo;
  double red[9] = {s[0], c->eng_vdwl + tp.eng_bond, c->virial[0] + tp.virial_bond[0], c->virial[1] + tp.virial_bond[1],
                   c->virial[2] + tp.virial_bond[2], c->virial[3] + tp.virial_bond[3], c->virial[4] + tp.virial_bond[4],
                   c->virial[5] + tp.virial_bond[5], tp.eng_bond};
  TRY(comm_allreduce_sum(c, red, 9));
  const double dof = 3.0 * c->natoms_global - 3.0;  // ComputeTemp::dof_compute, extra_dof = 3
  const double tfactor = dof > 0 ? c->mvv2e / (dof * c->boltz) : 0.0;
  const double t = red[0] * tfactor;
  const double vol = c->prd[0] * c->prd[1] * c->prd[2];
  out->step = c->step;
  out->natoms = c->natoms_global;
  out->temp = t;
  out->pe = red[1];
  out->ke = t * 0.5 * dof * c->boltz;  // Thermo::compute_ke
  for (int k = 0; k < 6; k++) out->virial[k] = red[2 + k];
  out->emol = red[8];
  out->press = (dof * c->boltz * t + red[2] + red[3] + red[4]) / 3.0 * (1.0 / vol) * c->nktv2p;
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// whole loop
// ---------------------------------------------------------------------------------------
extern "C" int b200md_setup(b200md_ctx *c) {
  NEED(c && c->neigh_configured, "b200md: neighbor not configured");
  NEED(c->pair_style != B200MD_PAIR_NONE, "b200md: no pair style configured");
  TRY(b200md_neigh_build(c));
  TRY(b200md_pair_compute(c, 1, 1, nullptr, nullptr));
  TRY(bond_compute(c, 1, 1));
  TRY(b200md_reverse_comm(c));
  if (c->lang.on) {   // Modify::setup -> FixLangevin::setup -> post_force on the set-up forces
    Langevin &L = c->lang;
    if (!L.window) L.begin = L.end = L.now = c->step;
    // the host sorted at its set-up (Atom::sort sets nextsort from the current timestep, atom.cpp:2252)
    if (L.sortfreq > 0) L.nextsort = (L.now / L.sortfreq) * L.sortfreq + L.sortfreq;
    TRY(langevin_post_force(c, false, 0.0));
  }
  return B200MD_OK;
}

// option trace: a timed event on stream `s` under a label (development aid: what the two streams overlap)
static void trace_mark(b200md_ctx *c, const char *label, cudaStream_t s) {
  if (!c->trace) return;
  cudaEvent_t e;
  cudaEventCreate(&e);
  cudaEventRecord(e, s);
  c->trace_ev.emplace_back(label, e);
}
static void trace_dump(b200md_ctx *c) {
  if (!c->trace || c->trace_ev.empty()) return;
  for (auto &le : c->trace_ev) {
    float ms = 0;
    cudaEventElapsedTime(&ms, c->trace_ev[0].second, le.second);
    fprintf(stderr, "[trace rank %d] %9.1f us  %s\n", c->rank, ms * 1e3, le.first.c_str());
    }
  for (auto &le : c->trace_ev) cudaEventDestroy(le.second);
  c->trace_ev.clear();
}

// Non-rebuild, non-thermo step on several ranks: the forward halo (and EAM's fp halo) runs on the
// comm stream while the tiles without ghost neighbors are evaluated; boundary tiles follow.
// Same kernels, same per-atom arithmetic and order as the unsplit launch -> identical results.
static bool can_overlap(const b200md_ctx *c) {
  return c->overlap && !c->all_self && c->n_tiles_int > 0 && !c->topo.molecular;
}
static int on_comm_stream(b200md_ctx *c, int (*fn)(b200md_ctx *)) {
  cudaStream_t saved = c->stream;
  c->stream = c->comm_stream;
  int r = fn(c);
  c->stream = saved;
  return r;
}
// Compute / halo overlap on several ranks (non-rebuild, non-tally steps).  Tiles whose atoms have no ghost neighbour
// ("interior", ~85-90 % at 2 M atoms per GPU) are ONE launch on the main stream.  The boundary tiles are a second
// launch on a high-priority stream that waits for the forward halo: their blocks are dispatched ahead of the interior
// kernel's remaining blocks as soon as the ghosts have landed, so the reverse pass (comm stream, FP64 RED into f / rho)
// is in flight while most of the interior work is still running, and neither kernel leaves a tail of idle SMs for
// the other to wait behind.  Same kernels, same per-atom arithmetic as the unsplit launch.
static int on_stream(b200md_ctx *c, cudaStream_t st, int (*fn)(b200md_ctx *, bool, int, const int *, int), bool ev, int pass,
                     const int *tiles, int ntl) {
  cudaStream_t saved = c->stream;
  c->stream = st;
  int r = fn(c, ev, pass, tiles, ntl);
  c->stream = saved;
  return r;
}
static int lj_pass(b200md_ctx *c, bool ev, int, const int *tiles, int ntl) { return pair_lj_launch(c, ev, tiles, ntl); }

// non-rebuild, non-tally step on several ranks; on return f is complete (reverse_comm included)
static int forces_overlapped(b200md_ctx *c) {
  cudaStream_t s = c->stream, cs = c->comm_stream, bs = c->bnd_stream;
  const bool half = c->list_kind == B200MD_LIST_HALF_NEWTON, eam = c->pair_style == B200MD_PAIR_EAM;
  const int nall = c->nlocal + c->nghost;
  const int *ti = c->tiles_int, *tb = c->tiles_bnd;
  const int ni = c->n_tiles_int, nb = c->n_tiles_bnd;
  trace_mark(c, "main: step start (x integrated)", s);
  CUDA_TRY(cudaEventRecord(c->ev_x, s));               // positions integrated
  CUDA_TRY(cudaStreamWaitEvent(cs, c->ev_x, 0));
  TRY(on_comm_stream(c, comm_forward));                // x halo on the comm stream
  trace_mark(c, "comm: x forward done", cs);
  CUDA_TRY(cudaEventRecord(c->ev_halo, cs));
  c->ev_valid = false;
  if (half) {   // Verlet::force_clear over owned + ghost rows
    CUDA_TRY(cudaMemset2DAsync(c->f, sizeof(double) * (size_t)c->nmax, 0, sizeof(double) * nall, 3, s));   // 3 rows, one launch
    if (eam) CUDA_TRY(cudaMemsetAsync(c->rho, 0, sizeof(double) * nall, s));
  }
  CUDA_TRY(cudaEventRecord(c->ev_clr, s));
  CUDA_TRY(cudaStreamWaitEvent(bs, c->ev_clr, 0));
  CUDA_TRY(cudaStreamWaitEvent(bs, c->ev_halo, 0));
  if (!eam) {
    TRY(lj_pass(c, false, 0, ti, ni));
    trace_mark(c, "main: lj interior done", s);
    TRY(on_stream(c, bs, lj_pass, false, 0, tb, nb));
    trace_mark(c, "bnd : lj boundary done", bs);
  } else {
    // pass 1 (density).  half: rho of both atoms of a pair, ghost rho returned, then the embedding; full: rho + fp fused
    const int p1 = half ? 2 : 0, p3 = half ? 3 : 1;
    TRY(pair_eam_launch(c, false, p1, ti, ni));
    trace_mark(c, "main: rho interior done", s);
    TRY(on_stream(c, bs, pair_eam_launch, false, p1, tb, nb));
    trace_mark(c, "bnd : rho boundary done", bs);
    CUDA_TRY(cudaEventRecord(c->ev_rb, bs));
    if (half) {
      CUDA_TRY(cudaStreamWaitEvent(cs, c->ev_rb, 0));
      TRY(on_comm_stream(c, comm_reverse_rho));
      trace_mark(c, "comm: rho reverse done", cs);
      CUDA_TRY(cudaEventRecord(c->ev_rr, cs));
      CUDA_TRY(cudaStreamWaitEvent(s, c->ev_rr, 0));
    }
    CUDA_TRY(cudaStreamWaitEvent(s, c->ev_rb, 0));
    if (half) { TRY(pair_eam_embed(c, false, nullptr)); trace_mark(c, "main: embed done", s); }
    CUDA_TRY(cudaEventRecord(c->ev_fp, s));              // fp of every owned atom is final
    CUDA_TRY(cudaStreamWaitEvent(cs, c->ev_fp, 0));
    TRY(on_comm_stream(c, pair_eam_fp_comm));            // fp halo on the comm stream
    trace_mark(c, "comm: fp forward done", cs);
    CUDA_TRY(cudaEventRecord(c->ev_fphalo, cs));
    TRY(pair_eam_launch(c, false, p3, ti, ni));
    trace_mark(c, "main: force interior done", s);
    CUDA_TRY(cudaStreamWaitEvent(bs, c->ev_fphalo, 0));
    TRY(on_stream(c, bs, pair_eam_launch, false, p3, tb, nb));
    trace_mark(c, "bnd : force boundary done", bs);
  }
  CUDA_TRY(cudaEventRecord(c->ev_bnd, bs));              // boundary tiles done: ghost rows of f complete
  if (half) {
    CUDA_TRY(cudaStreamWaitEvent(cs, c->ev_bnd, 0));
    TRY(on_comm_stream(c, comm_reverse_force));
    trace_mark(c, "comm: f reverse done", cs);
    CUDA_TRY(cudaEventRecord(c->ev_rev, cs));
    CUDA_TRY(cudaStreamWaitEvent(s, c->ev_rev, 0));
  }
  CUDA_TRY(cudaStreamWaitEvent(s, c->ev_bnd, 0));
  return B200MD_OK;
}

// ---------------------------------------------------------------------------------------
// option graph: small systems are launch-bound (7-13 launches of a few microseconds per step, and with `check yes` a
// host round trip for the rebuild decision on every step).  The steps between two list builds run as ONE graph
// launch: a WHILE conditional node whose body is a complete non-rebuild timestep
//     forward_comm -> pair -> reverse_comm -> final_integrate(n) + initial_integrate(n+1) with the distance test
// followed by a one-thread kernel that evaluates Neighbor::decide for step n+1 (src/neighbor.cpp:2309-2324: ago++,
// ago >= delay && ago % every == 0, then check_distance's flag) and sets the loop condition.  The loop leaves exactly
// on the step the reference rebuilds on — its first half-kick and drift already applied, like Verlet::run has them
// when it calls Neighbor::decide — and the host does the rebuild.  Reference counterpart to beat:
// verlet_kokkos.cpp:287-524 (resident loop, but a host decision per step).
// ---------------------------------------------------------------------------------------
__global__ void k_loop_cond(cudaGraphConditionalHandle handle, int *__restrict__ loop, int *__restrict__ flags, int every,
                            int delay, int check) {
  const int left = loop[0] - 1, ago = loop[1] + 1;
  const bool due = ago >= delay && ago % every == 0;
  const int rebuild = (due && (!check || flags[0])) ? 1 : 0;
  if (!rebuild) flags[0] = 0;     // the test of a step that is not due does not count (check_distance runs on due steps only)
  loop[0] = left; loop[1] = ago; loop[2] += 1; loop[3] = rebuild;
  cudaGraphSetConditional(handle, (left > 0 && !rebuild) ? 1u : 0u);
}

static void step_graph_free(b200md_ctx *c) {
  StepGraph &G = c->sg;
  if (G.exec) cudaGraphExecDestroy(G.exec);
  if (G.graph) cudaGraphDestroy(G.graph);
  G.exec = nullptr; G.graph = nullptr; G.built_for = -1;
}

static bool step_graph_usable(const b200md_ctx *c) {
  const StepGraph &G = c->sg;
  if (G.mode == 0 || c->nranks != 1 || c->lang.on || c->topo.molecular || !c->fuse_nve || c->phase_timing || c->trace) return false;
  if (G.mode < 0) return false;   // auto = off: measured on B200 the 32 000-atom steps are bound by the chain of small dependent
                                  // kernels on the GPU, not by host launches (in.lj 62 us/step host-driven, 68 us as a graph)
  return c->nlocal > 0;
}

static int step_graph_build(b200md_ctx *c, double dtv, double dtf) {
  StepGraph &G = c->sg;
  if (G.exec && G.built_for == c->nbuilds && G.dtv == dtv && G.dtf == dtf) return B200MD_OK;
  step_graph_free(c);
  if (!G.d_loop) {
    CUDA_TRY(cudaMalloc(&G.d_loop, 4 * sizeof(int)));
    CUDA_TRY(cudaMallocHost(&G.h_loop, 4 * sizeof(int)));
  }
  CUDA_TRY(cudaGraphCreate(&G.graph, 0));
  cudaGraphConditionalHandle handle;
  CUDA_TRY(cudaGraphConditionalHandleCreate(&handle, G.graph, 1, cudaGraphCondAssignDefault));
  cudaGraphNodeParams np = {cudaGraphNodeTypeConditional};
  np.conditional.handle = handle;
  np.conditional.type = cudaGraphCondTypeWhile;
  np.conditional.size = 1;
  cudaGraphNode_t node;
  CUDA_TRY(cudaGraphAddNode(&node, G.graph, nullptr, 0, &np));
  cudaGraph_t body = np.conditional.phGraph_out[0];
  const long long l0 = c->launches, p0 = c->pair_launches;
  CUDA_TRY(cudaStreamBeginCaptureToGraph(c->stream, body, nullptr, nullptr, 0, cudaStreamCaptureModeThreadLocal));
  int rc = comm_forward(c);
  if (rc == B200MD_OK) rc = b200md_pair_compute(c, 0, 0, nullptr, nullptr);
  if (rc == B200MD_OK) rc = comm_reverse_force(c);
  if (rc == B200MD_OK) rc = nve_final_initial(c, dtv, dtf, c->check != 0);
  if (rc == B200MD_OK) k_loop_cond<<<1, 1, 0, c->stream>>>(handle, G.d_loop, c->d_flags, c->every, c->delay, c->check);
  cudaGraph_t captured = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(c->stream, &captured);
  G.launches_per_body = (int)(c->launches - l0) + 1;
  G.pair_launches_per_body = (int)(c->pair_launches - p0);
  c->launches = l0; c->pair_launches = p0;
  if (rc != B200MD_OK) { step_graph_free(c); return rc; }
  if (ce != cudaSuccess) {
    step_graph_free(c);
    b200md_set_error(std::string("b200md: CUDA graph capture failed: ") + cudaGetErrorString(ce));
    return B200MD_ERR;
  }
  CUDA_TRY(cudaGraphInstantiate(&G.exec, G.graph, 0));
  G.built_for = c->nbuilds; G.dtv = dtv; G.dtf = dtf;
  return B200MD_OK;
}

// run up to `nbody` consecutive non-rebuild, non-tally steps; the first one has had its initial_integrate and its
// Neighbor::decide.  On return: *done steps are complete, the following step is pre-integrated and *rebuild holds its
// decision (c->ago already counts it).
static int step_graph_run(b200md_ctx *c, int nbody, double dtv, double dtf, int *done, int *rebuild) {
  StepGraph &G = c->sg;
  TRY(step_graph_build(c, dtv, dtf));
  G.h_loop[0] = nbody; G.h_loop[1] = c->ago; G.h_loop[2] = 0; G.h_loop[3] = 0;
  CUDA_TRY(cudaMemcpyAsync(G.d_loop, G.h_loop, 4 * sizeof(int), cudaMemcpyHostToDevice, c->stream));
  CUDA_TRY(cudaGraphLaunch(G.exec, c->stream));
  CUDA_TRY(cudaMemcpyAsync(G.h_loop, G.d_loop, 4 * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  *done = G.h_loop[2];
  *rebuild = G.h_loop[3];
  c->ago = G.h_loop[1];
  c->ev_valid = false;
  c->launches += (long long)*done * G.launches_per_body;
  c->pair_launches += (long long)*done * G.pair_launches_per_body;
  G.nlaunch++; G.nsteps += *done;
  if (*rebuild && c->check && c->ago == (c->every > c->delay ? c->every : c->delay)) c->ndanger++;
  return B200MD_OK;
}

struct PhaseTimer {
  b200md_ctx *c; bool on; size_t *used; std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> spans;
  cudaEvent_t get() {
    if (*used == c->ev_pool.size()) { cudaEvent_t e; cudaEventCreate(&e); c->ev_pool.push_back(e); }
    return c->ev_pool[(*used)++];
  }
  void begin(int cat) { if (!on) return; cudaEvent_t a = get(), b = get(); cudaEventRecord(a, c->stream); spans.push_back({cat, {a, b}}); }
  void end() { if (!on) return; cudaEventRecord(spans.back().second.second, c->stream); }
};

extern "C" int b200md_run(b200md_ctx *c, int nsteps, double dt, int thermo_every, b200md_thermo *thermo_out,
                          int max_thermo, int *n_thermo) {
  NEED(c && c->neigh_configured && c->nbuilds > 0, "b200md: run before setup");
  CUDA_TRY(cudaSetDevice(c->device));
  const double dtv = dt, dtf = 0.5 * dt * c->ftm2v;  // FixNVE::init (src/fix_nve.cpp:51-58)
  const long long first = c->step, last = c->step + nsteps;
  int nth = 0;
  size_t ev_used = 0;
  PhaseTimer T{c, c->phase_timing, &ev_used};
  const long long launches0 = c->launches, pl0 = c->pair_launches, nb0 = c->nbuilds;
  PhaseTimer TT{c, true, &ev_used};
  Langevin &LG = c->lang;
  if (LG.on) {
    TRY(langevin_set_dt(c, dt));   // FixLangevin::reset_dt
    if (!LG.window) { LG.begin = first; LG.end = last; LG.now = first; }
  }
  bool initial_done = false, predecided = false;
  int pre_flag = 0;
  TT.begin(0);
  for (int i = 0; i < nsteps; i++) {
    c->step++;
    LG.now++;
    // Integrate::ev_set: energy/virial only on thermo output steps
    const bool out_step = (thermo_every > 0 && (c->step - first) % thermo_every == 0) || c->step == last;
    const int ev = out_step ? 1 : 0;
    if (initial_done) initial_done = false;   // this step's first half-kick + drift ran fused with the previous step's second one
    else { T.begin(4); TRY(b200md_nve_initial(c, dtv, dtf)); T.end(); }
    int nflag = 0;
    if (predecided) { nflag = pre_flag; predecided = false; }   // Neighbor::decide of this step ran on the device (option graph)
    else TRY(b200md_neigh_decide(c, &nflag));
    if (!nflag && !ev && i + 1 < nsteps && step_graph_usable(c)) {
      // steps i .. i+m-1 are non-output steps that have a successor in this run: one graph launch
      int m = 0;
      for (int j = i; j + 1 < nsteps; j++) {
        const long long stepj = first + j + 1;
        if ((thermo_every > 0 && (stepj - first) % thermo_every == 0) || stepj == last) break;
        m++;
      }
      int done = 0;
      TRY(step_graph_run(c, m, dtv, dtf, &done, &pre_flag));
      c->step += done - 1;
      i += done - 1;
      initial_done = true;
      predecided = true;
      continue;
    }
    if (!nflag && !ev && can_overlap(c)) {
      T.begin(1); TRY(forces_overlapped(c)); T.end();   // includes reverse_comm on half lists
    } else {
      if (!nflag) { T.begin(3); TRY(comm_forward(c)); T.end(); }
      else {
        if (LG.on && LG.sortfreq > 0 && LG.now >= LG.nextsort) LG.resort_pending = true;
        T.begin(2); TRY(b200md_neigh_build(c)); T.end();
      }
      T.begin(1); TRY(b200md_pair_compute(c, ev, ev, nullptr, nullptr)); TRY(bond_compute(c, ev, ev)); T.end();
      T.begin(3); TRY(comm_reverse_force(c)); T.end();
    }
    // Modify::post_force (fix langevin) fused with FixNVE::final_integrate when the thermostat is on
    T.begin(4);
    if (LG.on) {
      // the thermostat's step counter (ramp, re-sort) moves at the top of the loop: nothing sits between this step's
      // final_integrate and the next step's initial_integrate on non-output steps
      const bool next = c->fuse_nve && !out_step && i + 1 < nsteps;
      TRY(langevin_post_force(c, true, dtf, next, dtv, next && decide_will_check(c) && c->nbuilds > 0));
      initial_done = next;
    } else if (c->fuse_nve && !out_step && i + 1 < nsteps) {
      // the distance check belongs to the NEXT step's initial_integrate; ago only moves in decide / build
      TRY(nve_final_initial(c, dtv, dtf, decide_will_check(c) && c->nbuilds > 0));
      initial_done = true;
    } else TRY(nve_final(c, dtf));
    T.end();
    if (out_step && thermo_out && nth < max_thermo) { TRY(b200md_thermo_compute(c, &thermo_out[nth])); nth++; }
  }
  TT.end();
  trace_mark(c, "main: run end", c->stream);
  if (c->topo.molecular) CUDA_TRY(cudaMemcpyAsync(c->h_flags + 4, c->d_flags + 4, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  CUDA_TRY(cudaStreamSynchronize(c->stream));
  if (c->topo.molecular && c->h_flags[4]) {   // bond kernel flags, at least once per run (bond.cu)
    const int fl = c->h_flags[4];
    CUDA_TRY(cudaMemsetAsync(c->d_flags + 4, 0, sizeof(int), c->stream));
    if (fl & 1) { b200md_set_error("Bond atoms missing"); return B200MD_ERR; }
    if (fl & 4) { b200md_set_error("Bad FENE bond"); return B200MD_ERR; }
  }
  if (n_thermo) *n_thermo = nth;
  trace_dump(c);
  float ms = 0;
  cudaEventElapsedTime(&ms, TT.spans[0].second.first, TT.spans[0].second.second);
  c->timing.total_ms = ms;
  c->timing.pair_ms = c->timing.neigh_ms = c->timing.comm_ms = c->timing.modify_ms = 0;
  for (auto &s : T.spans) {
    float m = 0;
    cudaEventElapsedTime(&m, s.second.first, s.second.second);
    if (s.first == 1) c->timing.pair_ms += m; else if (s.first == 2) c->timing.neigh_ms += m;
    else if (s.first == 3) c->timing.comm_ms += m; else c->timing.modify_ms += m;
  }
  c->timing.launches = c->launches - launches0;
  c->timing.pair_launches = c->pair_launches - pl0;
  c->timing.nbuilds = (int)(c->nbuilds - nb0);
  return B200MD_OK;
}

extern "C" int b200md_get_timing(b200md_ctx *c, b200md_timing *out) {
  NEED(c && out, "b200md: null argument");
  *out = c->timing;
  return B200MD_OK;
}
