#ifdef FIX_CLASS
// clang-format off
FixStyle(nve/b200,FixNVEB200);
// clang-format on
#else
#ifndef LMP_FIX_NVE_B200_H
#define LMP_FIX_NVE_B200_H

#include "b200_styles.h"
#include "fix_nve.h"

namespace LAMMPS_NS {

// fix nve/b200: FixNVE::init (dtv, dtf) is the reference's; the two half-kicks
// (src/fix_nve.cpp:64-141) run on the device.  Under run_style verlet/b200 the whole step loop
// is on the device and this fix only marks "NVE integration of group all".
class FixNVEB200 : public FixNVE {
 public:
  FixNVEB200(class LAMMPS *lmp, int narg, char **arg) : FixNVE(lmp, narg, arg) {}
  void initial_integrate(int) override;
  void final_integrate() override;
};

}    // namespace LAMMPS_NS
#endif
#endif
