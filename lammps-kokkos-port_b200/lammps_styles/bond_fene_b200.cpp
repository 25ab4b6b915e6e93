#include "bond_fene_b200.h"

#include "atom.h"
#include "error.h"

using namespace LAMMPS_NS;

int BondFENEB200::b200_push_bond(b200md_ctx *ctx)
{
  // 1-based per-type arrays filled by BondFENE::coeff
  return b200md_bond_fene_config(ctx, atom->nbondtypes, k, r0, epsilon, sigma);
}

void BondFENEB200::compute(int eflag, int vflag)
{
  ev_init(eflag, vflag);
  if (!B200Shared::ctx) error->all(FLERR, "Bond style fene/b200 requires run_style verlet/b200");
  if (b200md_bond_compute(B200Shared::ctx, eflag, vflag, &energy, virial)) error->one(FLERR, b200md_last_error());
}
