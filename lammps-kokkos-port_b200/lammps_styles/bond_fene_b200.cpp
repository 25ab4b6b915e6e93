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

int BondFENEB200::b200_peratom_bond(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n)
{
  ev_init(eflag, vflag);
  return b200_fetch_peratom(ctx, B200MD_PERATOM_BOND, eflag_atom, vflag_atom, eatom, vatom, rows, n);
}
