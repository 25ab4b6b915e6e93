#include "fix_nve_b200.h"

#include "error.h"

using namespace LAMMPS_NS;

void FixNVEB200::initial_integrate(int /*vflag*/)
{
  if (!B200Shared::ctx) error->all(FLERR, "Fix nve/b200 requires run_style verlet/b200");
  if (b200md_nve_initial(B200Shared::ctx, dtv, dtf)) error->one(FLERR, b200md_last_error());
}

void FixNVEB200::final_integrate()
{
  if (!B200Shared::ctx) error->all(FLERR, "Fix nve/b200 requires run_style verlet/b200");
  if (b200md_nve_final(B200Shared::ctx, dtf)) error->one(FLERR, b200md_last_error());
}
