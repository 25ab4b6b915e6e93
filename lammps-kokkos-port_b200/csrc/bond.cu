// bond.cu — molecular topology for bench/in.chain: tag->index map, FENE bonds.
//
// Reference:
//   NTopoBondAll::build   /root/reference/src/ntopo_bond_all.cpp:37-80  (bond list at every rebuild:
//                         atom->map(bond_atom) -> Domain::closest_image -> (i, j, type))
//   Domain::closest_image src/domain.cpp:1123-1151
//   BondFENE::compute     src/MOLECULE/bond_fene.cpp:53-129
//   Bond::ev_tally        src/bond.cpp:174-256 (newton_bond on: whole bond energy / virial once)
// B200 design: the topology is keyed by atom TAG (CSR tables replicated on every rank), so
// nothing per-atom has to be permuted by the bin sort or packed by the migration.  Each owned
// atom evaluates the bonds stored on it (newton_bond on: every bond is stored once); the
// partner's local index comes from a tag->index map rebuilt with the neighbor list (owned copy
// if there is one, else any ghost image), and the displacement is the minimum image of
// x_i - x_j — the same vector closest_image() selects, computed without the image chain.
// f_i, f_j are accumulated with FP64 RED (the partner may be another thread's atom).
#include "ctx.h"
#include "util.cuh"

#define BOND_THREADS 256
#define CUBEROOT2 1.25992104989487316476   // MathConst::MY_CUBEROOT2 (src/math_const.h)

void topo_free(b200md_ctx *c) {
  Topology &t = c->topo;
  void *p[] = {t.bond_off, t.bond_partner, t.bond_type, t.spec_off, t.spec_tag, t.spec_which, t.map, t.fene};
  for (void *q : p) if (q) cudaFree(q);
  t = Topology();
}

template <typename T>
static int up(T *&dst, const T *src, size_t n) {
  if (dst) { cudaFree(dst); dst = nullptr; }
  CUDA_TRY(cudaMalloc(&dst, sizeof(T) * (n ? n : 1)));
  if (n) CUDA_TRY(cudaMemcpy(dst, src, sizeof(T) * n, cudaMemcpyHostToDevice));
  return B200MD_OK;
}

int topo_set(b200md_ctx *c, long long natoms, const int *bond_off, const int *bond_partner, const int *bond_type,
             const int *spec_off, const int *spec_tag, const int *spec_which, const double special_lj[4],
             const double special_coul[4]) {
  Topology &t = c->topo;
  const size_t n1 = (size_t)natoms + 2;
  t.natoms = natoms;
  TRY(up(t.bond_off, bond_off, n1));
  TRY(up(t.bond_partner, bond_partner, (size_t)bond_off[natoms + 1]));
  TRY(up(t.bond_type, bond_type, (size_t)bond_off[natoms + 1]));
  TRY(up(t.spec_off, spec_off, n1));
  TRY(up(t.spec_tag, spec_tag, (size_t)spec_off[natoms + 1]));
  TRY(up(t.spec_which, spec_which, (size_t)spec_off[natoms + 1]));
  if (t.map) cudaFree(t.map);
  CUDA_TRY(cudaMalloc(&t.map, sizeof(int) * n1));
  // Neighbor::init (src/neighbor.cpp:430-452): 0 = both weights 0 (excluded), 1 = both 1 (plain
  // neighbor), 2 = anything else (index carries the special code)
  for (int k = 1; k <= 3; k++) {
    t.special_lj[k] = special_lj[k];
    if (special_lj[k] == 0.0 && special_coul[k] == 0.0) t.special_flag[k] = 0;
    else if (special_lj[k] == 1.0 && special_coul[k] == 1.0) t.special_flag[k] = 1;
    else t.special_flag[k] = 2;
    if (t.special_flag[k] == 2) {
      b200md_set_error("b200md: special_bonds weights other than 0 or 1 are not supported");
      return B200MD_ERR;
    }
  }
  t.molecular = true;
  return B200MD_OK;
}

int bond_fene_set(b200md_ctx *c, int nbondtypes, const double *k, const double *r0, const double *eps, const double *sigma) {
  std::vector<double> h((size_t)(nbondtypes + 1) * 4, 0.0);
  for (int t = 1; t <= nbondtypes; t++) { h[t * 4] = k[t]; h[t * 4 + 1] = r0[t]; h[t * 4 + 2] = eps[t]; h[t * 4 + 3] = sigma[t]; }
  c->topo.nbondtypes = nbondtypes;
  return up(c->topo.fene, h.data(), h.size());
}

// ---- tag -> local index ----------------------------------------------------------------------
__global__ void k_map_fill(int n, int *__restrict__ map) {
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) map[t] = -1;
}
__global__ void k_map_set(int first, int count, const int *__restrict__ tag, long long natoms, int *__restrict__ map) {
  int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= count) return;
  const int t = tag[first + k];
  if (t >= 1 && t <= natoms) map[t] = first + k;
}

int topo_build_map(b200md_ctx *c) {
  Topology &t = c->topo;
  if (!t.molecular) return B200MD_OK;
  const int n1 = (int)t.natoms + 1;
  k_map_fill<<<(n1 + 255) / 256, 256, 0, c->stream>>>(n1, t.map);
  if (c->nghost) k_map_set<<<(c->nghost + 255) / 256, 256, 0, c->stream>>>(c->nlocal, c->nghost, c->tag, t.natoms, t.map);
  if (c->nlocal) k_map_set<<<(c->nlocal + 255) / 256, 256, 0, c->stream>>>(0, c->nlocal, c->tag, t.natoms, t.map);  // owned copies win
  c->launches += 3;
  CUDA_TRY(cudaGetLastError());
  return B200MD_OK;
}

// ---- bond_style fene -------------------------------------------------------------------------
template <bool EV>
__global__ void __launch_bounds__(BOND_THREADS)
k_bond_fene(int nlocal, int nmax, const double4 *__restrict__ xq, const int *__restrict__ tag,
            const int *__restrict__ bond_off, const int *__restrict__ bond_partner, const int *__restrict__ bond_type,
            const int *__restrict__ map, const double *__restrict__ coef, double3 prd, int3 periodic,
            double *__restrict__ f, double *__restrict__ partials, int *__restrict__ flags) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double ev[7] = {0, 0, 0, 0, 0, 0, 0};
  if (i < nlocal) {
    const double4 qi = xq[i];
    const int t = tag[i];
    double fx = 0, fy = 0, fz = 0;
    for (int b = bond_off[t]; b < bond_off[t + 1]; b++) {
      const int j = map[bond_partner[b]];
      if (j < 0) { atomicOr(&flags[4], 1); continue; }      // "Bond atoms missing"
      const double4 qj = xq[j];
      double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
      // minimum image = the image Domain::closest_image picks
      if (periodic.x) { if (dx > 0.5 * prd.x) dx -= prd.x; else if (dx < -0.5 * prd.x) dx += prd.x; }
      if (periodic.y) { if (dy > 0.5 * prd.y) dy -= prd.y; else if (dy < -0.5 * prd.y) dy += prd.y; }
      if (periodic.z) { if (dz > 0.5 * prd.z) dz -= prd.z; else if (dz < -0.5 * prd.z) dz += prd.z; }
      const double *cf = coef + bond_type[b] * 4;
      const double k = cf[0], r0 = cf[1], eps = cf[2], sigma = cf[3];
      const double rsq = rsq_nofma(dx, dy, dz);
      const double r0sq = r0 * r0;
      double rlogarg = 1.0 - rsq / r0sq;
      if (rlogarg < 0.1) {
        atomicOr(&flags[4], rlogarg <= -3.0 ? 4 : 2);       // "Bad FENE bond" / "FENE bond too long"
        rlogarg = 0.1;
      }
      double fbond = -k / rlogarg;
      double sr6 = 0.0;
      const bool inner = rsq < CUBEROOT2 * sigma * sigma;
      if (inner) {
        const double sr2 = sigma * sigma / rsq;
        sr6 = sr2 * sr2 * sr2;
        fbond += 48.0 * eps * sr6 * (sr6 - 0.5) / rsq;
      }
      fx += dx * fbond; fy += dy * fbond; fz += dz * fbond;
      atomicAdd(&f[j], -dx * fbond);
      atomicAdd(&f[j + nmax], -dy * fbond);
      atomicAdd(&f[j + 2 * (size_t)nmax], -dz * fbond);
      if (EV) {
        double ebond = -0.5 * k * r0sq * log(rlogarg);
        if (inner) ebond += 4.0 * eps * sr6 * (sr6 - 1.0) + eps;
        ev[0] += ebond;
        ev[1] += dx * dx * fbond; ev[2] += dy * dy * fbond; ev[3] += dz * dz * fbond;
        ev[4] += dx * dy * fbond; ev[5] += dx * dz * fbond; ev[6] += dy * dz * fbond;
      }
    }
    if (bond_off[t + 1] > bond_off[t]) {
      atomicAdd(&f[i], fx); atomicAdd(&f[i + nmax], fy); atomicAdd(&f[i + 2 * (size_t)nmax], fz);
    }
  }
  if (EV) block_reduce_store<7>(ev, partials);
}

// Bond::compute after Pair::compute (src/verlet.cpp:318-324); adds into f
int bond_compute(b200md_ctx *c, int eflag, int vflag) {
  Topology &t = c->topo;
  if (!t.molecular || !t.fene) return B200MD_OK;
  const bool ev = eflag || vflag;
  const int nblocks = (c->nlocal + BOND_THREADS - 1) / BOND_THREADS;
  if (ev) { t.eng_bond = 0; for (int k = 0; k < 6; k++) t.virial_bond[k] = 0; }
  if (nblocks == 0) return B200MD_OK;
  const double3 prd = make_double3(c->prd[0], c->prd[1], c->prd[2]);
  const int3 per = make_int3(c->periodic[0], c->periodic[1], c->periodic[2]);
  // several ranks: a bond partner can be a ghost, its share goes home with reverse_comm
  // (newton_bond on).  The full-list pair kernels never touch ghost rows of f, so they are cleared
  // here (half lists clear owned + ghost rows before the pair kernel, Verlet::force_clear).
  if (c->list_kind == B200MD_LIST_FULL && !c->all_self && c->nghost)
    for (int d = 0; d < 3; d++)
      CUDA_TRY(cudaMemsetAsync(c->f + d * (size_t)c->nmax + c->nlocal, 0, sizeof(double) * c->nghost, c->stream));
  if (ev) {
    TRY(ensure_partials(c, nblocks));
    k_bond_fene<true><<<nblocks, BOND_THREADS, 0, c->stream>>>(c->nlocal, c->nmax, c->xq, c->tag, t.bond_off, t.bond_partner,
                                                             t.bond_type, t.map, t.fene, prd, per, c->f, c->partials, c->d_flags);
  } else {
    k_bond_fene<false><<<nblocks, BOND_THREADS, 0, c->stream>>>(c->nlocal, c->nmax, c->xq, c->tag, t.bond_off, t.bond_partner,
                                                              t.bond_type, t.map, t.fene, prd, per, c->f, c->partials, c->d_flags);
  }
  c->launches++;
  CUDA_TRY(cudaGetLastError());
  if (ev) {
    double out[7];
    TRY(reduce_partials(c, nblocks, 7, out));
    t.eng_bond = out[0];
    for (int k = 0; k < 6; k++) t.virial_bond[k] = out[1 + k];
    // the reduction synchronised the stream: the error flags of this evaluation are final
    int fl = 0;
    CUDA_TRY(cudaMemcpy(&fl, c->d_flags + 4, sizeof(int), cudaMemcpyDeviceToHost));
    if (fl) CUDA_TRY(cudaMemsetAsync(c->d_flags + 4, 0, sizeof(int), c->stream));   // read -> cleared
    if (fl & 1) { b200md_set_error("Bond atoms missing"); return B200MD_ERR; }
    if (fl & 4) { b200md_set_error("Bad FENE bond"); return B200MD_ERR; }
  }
  return B200MD_OK;
}
