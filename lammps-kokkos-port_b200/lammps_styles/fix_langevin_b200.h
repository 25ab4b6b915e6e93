#ifdef FIX_CLASS
// clang-format off
FixStyle(langevin/b200,FixLangevinB200);
// clang-format on
#else
#ifndef LMP_FIX_LANGEVIN_B200_H
#define LMP_FIX_LANGEVIN_B200_H

#include "b200_styles.h"
#include "fix_langevin.h"

namespace LAMMPS_NS {

// fix langevin/b200: the reference's argument parsing, fix_modify handling and per-type
// prefactors (FixLangevin constructor / init, src/fix_langevin.cpp:60-150, 286-298) are kept; the
// thermostat itself (post_force_templated, :576-713) runs inside the device step loop of
// run_style verlet/b200, drawing the SAME RanMars numbers the reference would draw (the device
// regenerates the serial stream by jump-ahead and indexes it by the host's local atom order, which
// verlet/b200 leaves in Atom::sort order like Verlet::setup does).
class FixLangevinB200 : public FixLangevin, public B200FixStyle {
 public:
  FixLangevinB200(class LAMMPS *lmp, int narg, char **arg);
  void init() override;
  void setup(int) override;
  void post_force(int) override;
  int b200_push_fix(b200md_ctx *ctx) override;
  int b200_begin_chunk(b200md_ctx *ctx) override;

 private:
  b200md_ctx *pushed_to;
};

}    // namespace LAMMPS_NS
#endif
#endif
