#ifdef PAIR_CLASS
// clang-format off
PairStyle(lj/cut/b200,PairLJCutB200);
// clang-format on
#else
#ifndef LMP_PAIR_LJ_CUT_B200_H
#define LMP_PAIR_LJ_CUT_B200_H

#include "b200_styles.h"
#include "pair_lj_cut.h"

namespace LAMMPS_NS {

// pair_style lj/cut/b200: settings/coeff/init_one/mixing/restart are the reference's
// (src/pair_lj_cut.cpp:422-556); the force loop (src/pair_lj_cut.cpp:69-137) runs in libb200md.so.
class PairLJCutB200 : public PairLJCut, public B200PairStyle {
 public:
  PairLJCutB200(class LAMMPS *lmp);
  void compute(int, int) override;
  int b200_push(b200md_ctx *ctx) override;
  int b200_peratom(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) override;
};

}    // namespace LAMMPS_NS
#endif
#endif
