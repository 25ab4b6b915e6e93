#ifdef NPAIR_CLASS
// clang-format off
// bit mask 0: Neighbor::choose_pair (src/neighbor.cpp:2096-2190) never selects these by itself (an unmodified
// NeighRequest has no flag for a new back end); VerletB200 invokes them on output steps
NPairStyle(full/bin/atomonly/b200, NPairFullBinB200, 0);
NPairStyle(half/bin/newton/b200, NPairHalfBinNewtonB200, 0);
// clang-format on
#else
#ifndef LMP_NPAIR_B200_H
#define LMP_NPAIR_B200_H

#include "b200_styles.h"
#include "npair.h"

#include <vector>

namespace LAMMPS_NS {

// NPair::build(NeighList *) (src/npair.h:35, invoked like src/neighbor.cpp:2470-2476) for host-side consumers of a
// neighbor list under run_style verlet/b200: compute coord/atom, compute rdf, a second pair style's list, ...
// The list is not built on the host: it is the device's resident list (b200md_neigh_export), re-indexed into the
// host's atom order — owned atoms by tag, ghosts by (tag, shifted position) against the ghosts Comm::borders made —
// and written into list->inum / ilist / numneigh / firstneigh (pages of MyPage<int>) like
// NPairFullBinAtomonly::build (src/npair_full_bin_atomonly.cpp:32-92) / NPairHalfBinAtomonlyNewton::build
// (src/npair_half_bin_atomonly_newton.cpp:33-113) do.  The device list holds every pair within the neighbor cutoff
// at the last rebuild, i.e. a superset of the pairs within the force cutoff now, which is what an occasional list of
// the reference is, too (Neighbor::build_one re-uses the bins of the last rebuild, src/neighbor.cpp:2528-2572).
class NPairB200 : public NPair {
 public:
  NPairB200(class LAMMPS *lmp, bool half_) : NPair(lmp), half(half_) {}
  void build(class NeighList *) override;

  // the device list of this output step in host indices (full adjacency of the owned atoms), shared by every list
  // built on the step
  struct Snapshot {
    bigint step = -1;
    long long nbuilds = -1;
    int nlocal = 0;
    std::vector<int> offset, neigh;    // CSR: neighbours of host atom i are neigh[offset[i] .. offset[i+1])
  };
  static Snapshot snap;
  static int refresh(class LAMMPS *lmp, b200md_ctx *ctx);    // 0 ok; else the message is in b200md_last_error / errmsg
  static const char *errmsg;

 protected:
  bool half;
};

class NPairFullBinB200 : public NPairB200 {
 public:
  NPairFullBinB200(class LAMMPS *lmp) : NPairB200(lmp, false) {}
};

class NPairHalfBinNewtonB200 : public NPairB200 {
 public:
  NPairHalfBinNewtonB200(class LAMMPS *lmp) : NPairB200(lmp, true) {}
};

}    // namespace LAMMPS_NS
#endif
#endif
