#ifdef BOND_CLASS
// clang-format off
BondStyle(fene/b200,BondFENEB200);
// clang-format on
#else
#ifndef LMP_BOND_FENE_B200_H
#define LMP_BOND_FENE_B200_H

#include "b200_styles.h"
#include "bond_fene.h"

namespace LAMMPS_NS {

// bond_style fene/b200: coeff / init_style / restart are the reference's
// (src/MOLECULE/bond_fene.cpp:146-230); the bond loop (:53-129) runs in libb200md.so on the
// implicit bond list the library derives from the tag-keyed topology (b200md_set_topology).
class BondFENEB200 : public BondFENE, public B200BondStyle {
 public:
  BondFENEB200(class LAMMPS *lmp) : BondFENE(lmp) {}
  void compute(int, int) override;
  int b200_push_bond(b200md_ctx *ctx) override;
  int b200_peratom_bond(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) override;
};

}    // namespace LAMMPS_NS
#endif
#endif
