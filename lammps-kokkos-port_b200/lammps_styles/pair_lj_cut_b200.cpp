#include "pair_lj_cut_b200.h"

#include "atom.h"
#include "error.h"

using namespace LAMMPS_NS;

PairLJCutB200::PairLJCutB200(LAMMPS *lmp) : PairLJCut(lmp)
{
  respa_enable = 0;
  no_virial_fdotr_compute = 1;    // the library returns the virial itself (per-pair tally)
}

int PairLJCutB200::b200_push(b200md_ctx *ctx)
{
  // tables filled by Pair::init -> init_one for every i<=j (src/pair.cpp:254-276); contiguous
  // (ntypes+1)^2 blocks (src/memory.h:144-160)
  return b200md_pair_lj_cut_config(ctx, atom->ntypes, &cutsq[0][0], &lj1[0][0], &lj2[0][0], &lj3[0][0],
                                   &lj4[0][0], &offset[0][0]);
}

void PairLJCutB200::compute(int eflag, int vflag)
{
  ev_init(eflag, vflag);
  if (!B200Shared::ctx) error->all(FLERR, "Pair style lj/cut/b200 requires run_style verlet/b200");
  if (b200md_pair_compute(B200Shared::ctx, eflag, vflag, &eng_vdwl, virial))
    error->one(FLERR, b200md_last_error());
}

int PairLJCutB200::b200_peratom(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n)
{
  ev_init(eflag, vflag);    // allocates / clears eatom, vatom (owned + ghost rows)
  return b200_fetch_peratom(ctx, B200MD_PERATOM_PAIR, eflag_atom, vflag_atom, eatom, vatom, rows, n);
}
