// fix langevin/b200 — see fix_langevin_b200.h.

#include "fix_langevin_b200.h"

#include "atom.h"
#include "comm.h"
#include "error.h"
#include "update.h"

#include <vector>

using namespace LAMMPS_NS;

FixLangevinB200::FixLangevinB200(LAMMPS *lmp, int narg, char **arg) :
    FixLangevin(lmp, narg, arg), pushed_to(nullptr)
{
}

/* ----------------------------------------------------------------------
   what the device thermostat covers: constant-style target temperature (with the
   Tstart -> Tstop ramp), `scale`, group all, per-type masses; tally no, zero no, gjf no
------------------------------------------------------------------------- */

void FixLangevinB200::init()
{
  FixLangevin::init();
  if (tstr) error->all(FLERR, "Fix langevin/b200 does not support a variable temperature");
  if (gjfflag || tallyflag || zeroflag || oflag || ascale != 0.0)
    error->all(FLERR, "Fix langevin/b200 supports the keywords scale only (gjf, tally, zero, omega, angmom: no)");
  if (temperature) error->all(FLERR, "Fix langevin/b200 does not support a temperature bias");
  if (igroup != 0) error->all(FLERR, "Fix langevin/b200 requires group all");
  if (atom->rmass) error->all(FLERR, "Fix langevin/b200 requires per-type masses");
  if (comm->nprocs > 1)
    error->all(FLERR, "Fix langevin/b200 runs on one rank (the random stream follows the local atom order)");
}

/* the set-up evaluation is thermostatted by b200md_setup (FixLangevin::setup -> post_force) */

void FixLangevinB200::setup(int /*vflag*/) {}

void FixLangevinB200::post_force(int /*vflag*/)
{
  error->all(FLERR, "Fix langevin/b200 requires run_style verlet/b200");
}

/* ----------------------------------------------------------------------
   called by verlet/b200 after the atoms are on the device, before the set-up evaluation.
   The generator is created once, like the reference's (constructor), and lives across runs.
------------------------------------------------------------------------- */

int FixLangevinB200::b200_push_fix(b200md_ctx *ctx)
{
  if (pushed_to != ctx) {
    std::vector<double> r(atom->ntypes + 1, 1.0);
    for (int i = 1; i <= atom->ntypes; i++) r[i] = ratio[i];
    if (b200md_fix_langevin(ctx, t_start, t_stop, t_period, seed + comm->me, r.data(), update->dt)) return 1;
    pushed_to = ctx;
  }
  // Verlet::run re-sorts the atoms every atom->sortfreq steps (src/verlet.cpp:285); the device re-ranks
  // its stream index at the same rebuild
  if (b200md_fix_langevin_sort(ctx, atom->sortfreq)) return 1;
  return b200_begin_chunk(ctx);
}

/* Update::beginstep / endstep of the enclosing run -> FixLangevin::compute_target's ramp */

int FixLangevinB200::b200_begin_chunk(b200md_ctx *ctx)
{
  bigint end = update->endstep;
  if (end < update->ntimestep) end = update->ntimestep;
  return b200md_fix_langevin_window(ctx, update->ntimestep, update->beginstep, end);
}
