// NPairFullBinB200 / NPairHalfBinNewtonB200 — see npair_b200.h.  Written against the reference's public class
// interfaces (NPair, NeighList, MyPage, Atom); no reference code.

#include "npair_b200.h"

#include "atom.h"
#include "comm.h"
#include "error.h"
#include "force.h"
#include "my_page.h"
#include "neigh_list.h"
#include "update.h"

#include <algorithm>
#include <cmath>
#include <unordered_map>

using namespace LAMMPS_NS;

NPairB200::Snapshot NPairB200::snap;
const char *NPairB200::errmsg = "";

/* ----------------------------------------------------------------------
   device list -> full adjacency of the owned atoms in host indices.
   Requires the host ghosts of this step (Comm::borders on the synced positions, VerletB200).
------------------------------------------------------------------------- */

int NPairB200::refresh(LAMMPS *lmp, b200md_ctx *ctx)
{
  Atom *atom = lmp->atom;
  long long nb = 0;
  b200md_neigh_stats(ctx, &nb, nullptr, nullptr, nullptr);
  if (snap.step == lmp->update->ntimestep && snap.nbuilds == nb) return 0;

  int nl = 0, ng = 0;
  if (b200md_get_nlocal(ctx, &nl, &ng)) return 1;
  if (nl != atom->nlocal) { errmsg = "verlet/b200: host and device disagree on the number of owned atoms"; return 2; }
  const int nall_host = atom->nlocal + atom->nghost;

  // device atoms (owned in device order, then ghosts): tag and position
  std::vector<double> dx(3 * (size_t) (nl + ng));
  std::vector<int> dtag(nl + ng), dtype(ng);
  if (nl && b200md_get_atoms(ctx, dx.data(), nullptr, nullptr, nullptr, dtag.data(), nullptr)) return 1;
  if (ng && b200md_get_ghosts(ctx, dx.data() + 3 * (size_t) nl, dtype.data(), dtag.data() + nl)) return 1;

  // host copies of every tag: the owned atom first, then its ghost images
  double **x = atom->x;
  tagint *tag = atom->tag;
  std::unordered_multimap<int, int> copies;
  copies.reserve(nall_host);
  for (int i = 0; i < nall_host; i++) copies.emplace((int) tag[i], i);
  auto nearest = [&](int t, const double *p, bool ghost_only, double &best) {
    int hit = -1;
    best = 1e300;
    auto range = copies.equal_range(t);
    for (auto it = range.first; it != range.second; ++it) {
      const int h = it->second;
      if (ghost_only && h < atom->nlocal) continue;
      const double d = std::fabs(x[h][0] - p[0]) + std::fabs(x[h][1] - p[1]) + std::fabs(x[h][2] - p[2]);
      if (d < best) { best = d; hit = h; }
    }
    return hit;
  };
  // A device ghost was selected at the last rebuild, the host ghosts just now: one that has since drifted out of the
  // send slab has no host counterpart.  It is then farther than the neighbor cutoff from the sub-box, i.e. beyond the
  // force cutoff of every owned atom, and its pairs can be left out of a list whose consumers cut at the force cutoff.
  std::vector<int> dev2host(nl + ng);
  for (int k = 0; k < nl + ng; k++) {
    double d;
    int h = nearest(dtag[k], &dx[3 * (size_t) k], k >= nl, d);
    // same arithmetic on both sides (x + pbc*prd, AtomVec::pack_border): matching positions agree to the last bit
    if (h >= 0 && d > 1e-9) h = -1;
    if (k < nl && (h < 0 || h >= atom->nlocal)) { errmsg = "verlet/b200: an owned device atom has no host counterpart"; return 2; }
    dev2host[k] = h;
  }

  // device list (local indices), either kind
  std::vector<int> nn(nl);
  long long total = 0;
  if (b200md_neigh_export(ctx, 0, nn.data(), nullptr, &total)) return 1;
  std::vector<int> flat((size_t) std::max<long long>(total, 1));
  if (total && b200md_neigh_export(ctx, 0, nn.data(), flat.data(), &total)) return 1;
  const bool dev_half = lmp->force->newton_pair != 0;    // VerletB200: newton on -> half/newton list on the device

  std::vector<std::vector<int>> adj(nl);
  size_t p = 0;
  for (int i = 0; i < nl; i++) {
    const int hi = dev2host[i];
    for (int m = 0; m < nn[i]; m++, p++) {
      const int j = flat[p], hj = dev2host[j];
      if (hj < 0) continue;
      adj[hi].push_back(hj);
      if (!dev_half) continue;
      // a half list stores the pair once: give the partner its entry, too
      if (j < nl) adj[hj].push_back(hi);
      else {
        // the partner is a ghost image: its owner sees the image of i displaced the other way
        int owner = -1;
        auto range = copies.equal_range(dtag[j]);
        for (auto it = range.first; it != range.second; ++it)
          if (it->second < atom->nlocal) owner = it->second;
        if (owner < 0) { errmsg = "verlet/b200: list publication with a half list needs all owners on this rank"; return 2; }
        const double want[3] = {dx[3 * (size_t) i] - (dx[3 * (size_t) j] - x[owner][0]),
                                dx[3 * (size_t) i + 1] - (dx[3 * (size_t) j + 1] - x[owner][1]),
                                dx[3 * (size_t) i + 2] - (dx[3 * (size_t) j + 2] - x[owner][2])};
        double d;
        const int hk = nearest(dtag[i], want, false, d);
        if (hk >= 0 && d < 1e-6) adj[owner].push_back(hk);
      }
    }
  }
  snap.offset.assign(nl + 1, 0);
  for (int i = 0; i < nl; i++) snap.offset[i + 1] = snap.offset[i] + (int) adj[i].size();
  snap.neigh.resize(snap.offset[nl]);
  for (int i = 0; i < nl; i++) {
    std::sort(adj[i].begin(), adj[i].end());
    std::copy(adj[i].begin(), adj[i].end(), snap.neigh.begin() + snap.offset[i]);
  }
  snap.nlocal = nl;
  snap.step = lmp->update->ntimestep;
  snap.nbuilds = nb;
  return 0;
}

/* ----------------------------------------------------------------------
   fill the host list.  Full: every neighbour.  Half: each pair once — owned partner j > i; ghost partner with
   newton on only if it is not below i in (z,y,x) order (the antisymmetric tie-break of
   npair_half_bin_atomonly_newton.cpp:61-69 applied to every ghost, so exactly one of the two periodic images of a
   pair survives), with newton off always (npair_half_bin_newtoff.cpp: owned-ghost pairs are kept by both owners).
------------------------------------------------------------------------- */

void NPairB200::build(NeighList *list)
{
  if (B200Shared::ctx == nullptr) error->all(FLERR, "NPair b200 styles need run_style verlet/b200");
  const int rc = refresh(lmp, B200Shared::ctx);
  if (rc == 1) error->all(FLERR, b200md_last_error());
  if (rc == 2) error->all(FLERR, errmsg);

  const int nlocal = atom->nlocal;
  double **x = atom->x;
  const bool newton = force->newton_pair != 0;
  list->grow(nlocal, nlocal + atom->nghost);
  int *ilist = list->ilist;
  int *numneigh = list->numneigh;
  int **firstneigh = list->firstneigh;
  MyPage<int> *ipage = list->ipage;
  ipage->reset();

  for (int i = 0; i < nlocal; i++) {
    int n = 0;
    int *neighptr = ipage->vget();
    for (int p = snap.offset[i]; p < snap.offset[i + 1]; p++) {
      const int j = snap.neigh[p];
      if (half) {
        if (j < nlocal) {
          if (j <= i) continue;
        } else if (newton) {
          if (x[j][2] < x[i][2]) continue;
          if (x[j][2] == x[i][2]) {
            if (x[j][1] < x[i][1]) continue;
            if (x[j][1] == x[i][1] && x[j][0] < x[i][0]) continue;
          }
        }
      }
      neighptr[n++] = j;
    }
    ilist[i] = i;
    firstneigh[i] = neighptr;
    numneigh[i] = n;
    ipage->vgot(n);
    if (ipage->status()) error->one(FLERR, "Neighbor list overflow, boost neigh_modify one");
  }
  list->inum = nlocal;
  list->gnum = 0;
}
