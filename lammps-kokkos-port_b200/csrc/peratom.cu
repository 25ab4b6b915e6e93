// peratom.cu — per-atom energy and virial on output steps.
//
// Reference:
//   Pair::ev_tally           /root/reference/src/pair.cpp:1036-1040 (eatom), :1078-1097 (vatom):
//                            half of evdwl and of v = del (x) del * fpair to each atom of a pair
//   PairEAM::compute         src/MANYBODY/pair_eam.cpp:245  eatom[i] += scale_ii * F(rho_i)
//   Bond::ev_tally           src/bond.cpp:186-192 (eatom), :234-251 (vatom)
//   ComputePEAtom            src/compute_pe_atom.cpp:92-170   pair + bond, reverse_comm of ghosts
//   ComputeStressAtom        src/compute_stress_atom.cpp:140-240  same for the 6 virial components
// B200 design: these tallies are wanted on dump / thermo-with-per-atom-compute steps only, so they
// do not ride in the per-step force kernels (whose register budget is what makes them fast,
// pair_lj.cu).  A separate pass walks the current neighbor list — either kind — at the current
// positions, one thread per owned atom, and writes seven SoA columns pa[k][nmax]
// (k = 0 energy, 1..6 xx yy zz xy xz yz).  Full list: every pair is seen from both atoms, each
// keeps its half, no atomics, no communication.  Half list / bonds: the partner's half goes out
// with FP64 RED and ghost rows travel back through the same reverse path the forces use
// (comm_reverse_scalar per column).  EAM reuses rho / fp of the last force evaluation, which is
// why the entry point requires the forces to be current.
#include "ctx.h"
#include "util.cuh"

#define PA_THREADS 128
#define CUBEROOT2 1.25992104989487316476   // MathConst::MY_CUBEROOT2 (src/math_const.h)

void peratom_free(b200md_ctx *c) {
  if (c->peratom) cudaFree(c->peratom);
  c->peratom = nullptr;
  c->peratom_cap = 0;
}

// slot k of atom i in the warp-tile transposed list written for F force lanes (ctx.h)
__device__ __forceinline__ const int *list_base(const int *neigh, int pitch, int F, int i) {
  const int A = 32 / F;
  return neigh + (size_t)(i / A) * pitch * A + (i % A) * F;
}
__device__ __forceinline__ int list_at(const int *base, int F, int k) { return base[(k / F) * 32 + (k % F)]; }

struct Tally {
  double e, v[6];
  __device__ void clear() { e = 0; for (int k = 0; k < 6; k++) v[k] = 0; }
};

// half of the pair's energy / virial to i (registers) and, when the pair is listed once, to j (RED)
template <bool BOTH>
__device__ __forceinline__ void tally_pair(Tally &t, double *__restrict__ pa, size_t nmax, int j, bool do_e, bool do_v,
                                           double e, double fpair, double dx, double dy, double dz) {
  if (do_e) {
    t.e += 0.5 * e;
    if (BOTH) atomicAdd(&pa[j], 0.5 * e);
  }
  if (do_v) {
    const double v[6] = {dx * dx * fpair, dy * dy * fpair, dz * dz * fpair, dx * dy * fpair, dx * dz * fpair, dy * dz * fpair};
#pragma unroll
    for (int k = 0; k < 6; k++) {
      t.v[k] += 0.5 * v[k];
      if (BOTH) atomicAdd(&pa[(size_t)(k + 1) * nmax + j], 0.5 * v[k]);
    }
  }
}
__device__ __forceinline__ void tally_store(const Tally &t, double *__restrict__ pa, size_t nmax, int i, bool do_e, bool do_v) {
  // RED in every case: other threads may be adding their half of a pair or bond to row i
  if (do_e) atomicAdd(&pa[i], t.e);
  if (do_v) {
#pragma unroll
    for (int k = 0; k < 6; k++) atomicAdd(&pa[(size_t)(k + 1) * nmax + i], t.v[k]);
  }
}

// ---- pair_style lj/cut (pair_lj_cut.cpp:69-137) ------------------------------------------------
template <bool HALF>
__global__ void __launch_bounds__(PA_THREADS)
k_lj_atom(int nlocal, int nmax, const double4 *__restrict__ xq, const int *__restrict__ neigh,
          const int *__restrict__ numneigh, int pitch, int F, int cl, const double *__restrict__ ljtab, int ntypes,
          double *__restrict__ pa, int do_e, int do_v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const int nt1 = ntypes + 1;
  const double4 qi = xq[i];
  const int itype = xq_type(qi);
  // cluster-pair list (cl): the list unit is the cluster i / CL, atom i owns the entries with bit i % CL of the mask
  const int unit = cl ? i / CL : i;
  const int n = numneigh[unit];
  const int *base = list_base(neigh, pitch, F, unit);
  Tally t; t.clear();
  for (int k = 0; k < n; k++) {
    int j = list_at(base, F, k);
    if (cl) {
      if (!(((unsigned)j >> (CL_SHIFT + i % CL)) & 1u)) continue;
      j &= CL_JMASK;
    }
    const double4 qj = xq[j];
    const double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
    const double rsq = rsq_nofma(dx, dy, dz);
    const double *c = ljtab + (size_t)(itype * nt1 + xq_type(qj)) * 8;   // cutsq lj1 lj2 lj3 lj4 offset
    if (rsq < c[0]) {
      const double r2inv = 1.0 / rsq;
      const double r6inv = r2inv * r2inv * r2inv;
      const double fpair = r6inv * (c[1] * r6inv - c[2]) * r2inv;
      const double evdwl = r6inv * (c[3] * r6inv - c[4]) - c[5];
      tally_pair<HALF>(t, pa, nmax, j, do_e, do_v, evdwl, fpair, dx, dy, dz);
    }
  }
  tally_store(t, pa, nmax, i, do_e, do_v);
}

// ---- pair_style eam / eam/alloy / eam/fs (pair_eam.cpp:231-247 embedding, :257-325 pair part) ---
template <bool HALF>
__global__ void __launch_bounds__(PA_THREADS)
k_eam_atom(int nlocal, int nmax, const double4 *__restrict__ xq, const int *__restrict__ neigh,
           const int *__restrict__ numneigh, int pitch, int F, int cl, EamParams P, const double *__restrict__ rho,
           const double *__restrict__ fp, double *__restrict__ pa, int do_e, int do_v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const int nt1 = P.ntypes + 1;
  const double4 qi = xq[i];
  const int itype = xq_type(qi);
  const double fpi = fp[i];
  // cluster-pair list (cl): the list unit is the cluster i / CL, atom i owns the entries with bit i % CL of the mask
  const int unit = cl ? i / CL : i;
  const int n = numneigh[unit];
  const int *base = list_base(neigh, pitch, F, unit);
  const size_t rows = (size_t)P.nr + 1;
  Tally t; t.clear();
  if (do_e) {   // embedding energy of atom i
    const double r = rho[i];
    double p = r * P.rdrho + 1.0;
    int m = (int)p;
    m = max(1, min(m, P.nrho - 1));
    p -= m;
    p = fmin(p, 1.0);
    const double *cf = P.frho + ((size_t)P.type2frho[itype] * (P.nrho + 1) + m) * 8;
    double phi = ((cf[3] * p + cf[4]) * p + cf[5]) * p + cf[6];
    if (r > P.rhomax) phi += fpi * (r - P.rhomax);
    t.e += phi * P.scale[itype * nt1 + itype];
  }
  for (int k = 0; k < n; k++) {
    int j = list_at(base, F, k);
    if (cl) {
      if (!(((unsigned)j >> (CL_SHIFT + i % CL)) & 1u)) continue;
      j &= CL_JMASK;
    }
    const double4 qj = xq[j];
    const double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
    const double rsq = rsq_nofma(dx, dy, dz);
    if (rsq < P.cutforcesq) {
      const int jtype = xq_type(qj);
      const double r = sqrt(rsq);
      double p = r * P.rdr + 1.0;
      int m = (int)p;
      m = min(m, P.nr - 1);
      p -= m;
      p = fmin(p, 1.0);
      const double *a = P.rhoA + ((size_t)P.type2rhor[itype * nt1 + jtype] * rows + m) * 4;
      const double *b = P.rhoA + ((size_t)P.type2rhor[jtype * nt1 + itype] * rows + m) * 4;
      const double *z = P.z2rV + ((size_t)P.type2z2r[itype * nt1 + jtype] * rows + m) * 4;
      const double rhoip = ((3.0 * a[0] * p + 2.0 * a[1]) * p + a[2]) * P.rdr;
      const double rhojp = ((3.0 * b[0] * p + 2.0 * b[1]) * p + b[2]) * P.rdr;
      const double z2p = ((3.0 * z[0] * p + 2.0 * z[1]) * p + z[2]) * P.rdr;
      const double z2 = ((z[0] * p + z[1]) * p + z[2]) * p + z[3];
      const double recip = 1.0 / r;
      const double phi = z2 * recip;
      const double phip = z2p * recip - phi * recip;
      const double psip = fpi * rhojp + fp[j] * rhoip + phip;
      const double scl = P.scale[itype * nt1 + jtype];
      tally_pair<HALF>(t, pa, nmax, j, do_e, do_v, scl * phi, -scl * psip * recip, dx, dy, dz);
    }
  }
  tally_store(t, pa, nmax, i, do_e, do_v);
}

// ---- bond_style fene (bond_fene.cpp:53-129), same bond walk as k_bond_fene (bond.cu) -----------
__global__ void __launch_bounds__(PA_THREADS)
k_fene_atom(int nlocal, int nmax, const double4 *__restrict__ xq, const int *__restrict__ tag,
            const int *__restrict__ bond_off, const int *__restrict__ bond_partner, const int *__restrict__ bond_type,
            const int *__restrict__ map, const double *__restrict__ coef, double3 prd, int3 periodic,
            double *__restrict__ pa, int do_e, int do_v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nlocal) return;
  const double4 qi = xq[i];
  const int tg = tag[i];
  if (bond_off[tg + 1] == bond_off[tg]) return;
  Tally t; t.clear();
  for (int b = bond_off[tg]; b < bond_off[tg + 1]; b++) {
    const int j = map[bond_partner[b]];
    if (j < 0) continue;   // reported by the force evaluation ("Bond atoms missing")
    const double4 qj = xq[j];
    double dx = qi.x - qj.x, dy = qi.y - qj.y, dz = qi.z - qj.z;
    if (periodic.x) { if (dx > 0.5 * prd.x) dx -= prd.x; else if (dx < -0.5 * prd.x) dx += prd.x; }
    if (periodic.y) { if (dy > 0.5 * prd.y) dy -= prd.y; else if (dy < -0.5 * prd.y) dy += prd.y; }
    if (periodic.z) { if (dz > 0.5 * prd.z) dz -= prd.z; else if (dz < -0.5 * prd.z) dz += prd.z; }
    const double *cf = coef + bond_type[b] * 4;
    const double k = cf[0], r0 = cf[1], eps = cf[2], sigma = cf[3];
    const double rsq = rsq_nofma(dx, dy, dz);
    const double r0sq = r0 * r0;
    double rlogarg = 1.0 - rsq / r0sq;
    if (rlogarg < 0.1) rlogarg = 0.1;
    double fbond = -k / rlogarg;
    double ebond = -0.5 * k * r0sq * log(rlogarg);
    if (rsq < CUBEROOT2 * sigma * sigma) {
      const double sr2 = sigma * sigma / rsq;
      const double sr6 = sr2 * sr2 * sr2;
      fbond += 48.0 * eps * sr6 * (sr6 - 0.5) / rsq;
      ebond += 4.0 * eps * sr6 * (sr6 - 1.0) + eps;
    }
    tally_pair<true>(t, pa, nmax, j, do_e, do_v, ebond, fbond, dx, dy, dz);
  }
  tally_store(t, pa, nmax, i, do_e, do_v);
}

// pa[0] (energy) and pa[1..6] (virial) of owned atoms for the current positions / list / rho / fp
int peratom_compute(b200md_ctx *c, int eflag_atom, int vflag_atom, int which) {
  if (c->peratom_cap < c->nmax) {
    peratom_free(c);
    CUDA_TRY(cudaMalloc(&c->peratom, sizeof(double) * 7 * (size_t)c->nmax));
    c->peratom_cap = c->nmax;
  }
  // columns are addressed with the context's nmax stride like f / v
  const int nmax = c->nmax, nall = c->nlocal + c->nghost;
  double *pa = c->peratom;
  for (int k = 0; k < 7; k++) CUDA_TRY(cudaMemsetAsync(pa + (size_t)k * nmax, 0, sizeof(double) * nall, c->stream));
  const int nblocks = (c->nlocal + PA_THREADS - 1) / PA_THREADS;
  const bool half = c->list_kind == B200MD_LIST_HALF_NEWTON;
  bool ghost_rows = false;
  if (nblocks && (which & B200MD_PERATOM_PAIR)) {
    if (c->pair_style == B200MD_PAIR_LJ_CUT) {
      if (half) k_lj_atom<true><<<nblocks, PA_THREADS, 0, c->stream>>>(c->nlocal, nmax, c->xq, c->neigh, c->numneigh, c->pitch, c->tpa, c->cluster ? 1 : 0,
                                                                      c->ljtab, c->ntypes, pa, eflag_atom, vflag_atom);
      else k_lj_atom<false><<<nblocks, PA_THREADS, 0, c->stream>>>(c->nlocal, nmax, c->xq, c->neigh, c->numneigh, c->pitch, c->tpa, c->cluster ? 1 : 0,
                                                                  c->ljtab, c->ntypes, pa, eflag_atom, vflag_atom);
    } else if (c->pair_style == B200MD_PAIR_EAM) {
      const EamParams P = pair_eam_params(c);
      if (half) k_eam_atom<true><<<nblocks, PA_THREADS, 0, c->stream>>>(c->nlocal, nmax, c->xq, c->neigh, c->numneigh, c->pitch, c->tpa, c->cluster ? 1 : 0,
                                                                       P, c->rho, c->fp, pa, eflag_atom, vflag_atom);
      else k_eam_atom<false><<<nblocks, PA_THREADS, 0, c->stream>>>(c->nlocal, nmax, c->xq, c->neigh, c->numneigh, c->pitch, c->tpa, c->cluster ? 1 : 0,
                                                                   P, c->rho, c->fp, pa, eflag_atom, vflag_atom);
    } else {
      b200md_set_error("All pair coeffs are not set");
      return B200MD_ERR;
    }
    c->launches++;
    ghost_rows = half;
  }
  const Topology &t = c->topo;
  if (nblocks && (which & B200MD_PERATOM_BOND) && t.molecular && t.fene) {
    const double3 prd = make_double3(c->prd[0], c->prd[1], c->prd[2]);
    const int3 per = make_int3(c->periodic[0], c->periodic[1], c->periodic[2]);
    k_fene_atom<<<nblocks, PA_THREADS, 0, c->stream>>>(c->nlocal, nmax, c->xq, c->tag, t.bond_off, t.bond_partner, t.bond_type,
                                                      t.map, t.fene, prd, per, pa, eflag_atom, vflag_atom);
    c->launches++;
    ghost_rows = true;
  }
  CUDA_TRY(cudaGetLastError());
  // ComputePEAtom / ComputeStressAtom::reverse_comm: ghost rows back to their owners
  if (ghost_rows) {
    if (eflag_atom) TRY(comm_reverse_column(c, pa));
    if (vflag_atom) for (int k = 1; k < 7; k++) TRY(comm_reverse_column(c, pa + (size_t)k * nmax));
  }
  return B200MD_OK;
}
