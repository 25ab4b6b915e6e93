#ifdef PAIR_CLASS
// clang-format off
PairStyle(eam/b200,PairEAMB200);
PairStyle(eam/alloy/b200,PairEAMAlloyB200);
PairStyle(eam/fs/b200,PairEAMFSB200);
// clang-format on
#else
#ifndef LMP_PAIR_EAM_B200_H
#define LMP_PAIR_EAM_B200_H

#include "b200_styles.h"
#include "pair_eam.h"
#include "pair_eam_alloy.h"
#include "pair_eam_fs.h"

namespace LAMMPS_NS {

// pair_style eam/b200, eam/alloy/b200 and eam/fs/b200: potential-file reading and spline construction are the
// reference's (src/MANYBODY/pair_eam.cpp:370-797, pair_eam_alloy.cpp:43-337); the three passes
// of PairEAM::compute (pair_eam.cpp:142-328) run in libb200md.so.
class PairEAMB200 : public PairEAM, public B200PairStyle {
 public:
  PairEAMB200(class LAMMPS *lmp) : PairEAM(lmp) { no_virial_fdotr_compute = 1; }
  void compute(int, int) override;
  int b200_push(b200md_ctx *ctx) override;
  int b200_peratom(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) override;
};

class PairEAMAlloyB200 : public PairEAMAlloy, public B200PairStyle {
 public:
  PairEAMAlloyB200(class LAMMPS *lmp) : PairEAM(lmp), PairEAMAlloy(lmp) { no_virial_fdotr_compute = 1; }
  void compute(int, int) override;
  int b200_push(b200md_ctx *ctx) override;
  int b200_peratom(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) override;
};

class PairEAMFSB200 : public PairEAMFS, public B200PairStyle {
 public:
  PairEAMFSB200(class LAMMPS *lmp) : PairEAM(lmp), PairEAMFS(lmp) { no_virial_fdotr_compute = 1; }
  void compute(int, int) override;
  int b200_push(b200md_ctx *ctx) override;
  int b200_peratom(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) override;
};

}    // namespace LAMMPS_NS
#endif
#endif
