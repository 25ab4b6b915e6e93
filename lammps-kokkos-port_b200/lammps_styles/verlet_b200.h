#ifdef INTEGRATE_CLASS
// clang-format off
IntegrateStyle(verlet/b200,VerletB200);
// clang-format on
#else
#ifndef LMP_VERLET_B200_H
#define LMP_VERLET_B200_H

#include "b200_styles.h"
#include "verlet.h"

#include <string>
#include <utility>
#include <vector>

namespace LAMMPS_NS {

// run_style verlet/b200: Verlet::setup / Verlet::run (src/verlet.cpp:93-162, 229-360) with the
// whole step loop on the device.  Host arrays are refreshed on output steps only, the way
// verlet/kk syncs them (src/KOKKOS/verlet_kokkos.cpp:287,516).
class VerletB200 : public Verlet {
 public:
  VerletB200(class LAMMPS *lmp, int narg, char **arg);
  ~VerletB200() override;
  void init() override;
  void setup(int flag) override;
  void setup_minimal(int flag) override;
  void run(int n) override;
  void cleanup() override;

 private:
  b200md_ctx *ctx;
  bool host_stale;
  long long builds_at_setup, danger_at_setup;
  std::vector<std::pair<std::string, std::string>> options;    // run_style verlet/b200 keyword value ...
  std::vector<std::pair<int, int>> mask_by_tag;    // several ranks: only when group masks are not uniform
  std::vector<std::pair<int, int>> host_index;     // one rank: (tag, host index), sorted; host order is kept
  std::vector<double> hx, hv, hf;                   // staging for the order-preserving sync
  std::vector<int> htype, htag, himage;

  void fail();
  void check_supported();
  void device_setup();
  void push_topology();
  void publish_tallies();
  void publish_peratom();
  void publish_lists();
  void host_borders();
  bigint ghosts_step;    // timestep the host ghosts (Comm::borders on the synced positions) belong to
  void sync_host();
};

}    // namespace LAMMPS_NS
#endif
#endif
