#include "pair_eam_b200.h"

#include "atom.h"
#include "error.h"

using namespace LAMMPS_NS;

// the reference's [ntable][n+1][7] spline blocks and 1-based type maps are passed as they are
#define B200_EAM_PUSH()                                                                              \
  b200md_pair_eam_config(ctx, atom->ntypes, nrho, nr, rdr, rdrho, rhomax, cutforcesq, nfrho,         \
                         &frho_spline[0][0][0], nrhor, &rhor_spline[0][0][0], nz2r,                  \
                         &z2r_spline[0][0][0], type2frho, &type2rhor[0][0], &type2z2r[0][0],         \
                         &scale[0][0])

#define B200_EAM_COMPUTE(name)                                                                       \
  ev_init(eflag, vflag);                                                                             \
  if (!B200Shared::ctx) error->all(FLERR, "Pair style " name " requires run_style verlet/b200");     \
  if (b200md_pair_compute(B200Shared::ctx, eflag, vflag, &eng_vdwl, virial))                         \
    error->one(FLERR, b200md_last_error())

#define B200_EAM_PERATOM()                                                                          \
  ev_init(eflag, vflag);                                                                             \
  return b200_fetch_peratom(ctx, B200MD_PERATOM_PAIR, eflag_atom, vflag_atom, eatom, vatom, rows, n)

int PairEAMB200::b200_peratom(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) { B200_EAM_PERATOM(); }
int PairEAMAlloyB200::b200_peratom(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) { B200_EAM_PERATOM(); }
int PairEAMFSB200::b200_peratom(b200md_ctx *ctx, int eflag, int vflag, const int *rows, int n) { B200_EAM_PERATOM(); }
int PairEAMB200::b200_push(b200md_ctx *ctx) { return B200_EAM_PUSH(); }
void PairEAMB200::compute(int eflag, int vflag) { B200_EAM_COMPUTE("eam/b200"); }
int PairEAMAlloyB200::b200_push(b200md_ctx *ctx) { return B200_EAM_PUSH(); }
void PairEAMAlloyB200::compute(int eflag, int vflag) { B200_EAM_COMPUTE("eam/alloy/b200"); }
int PairEAMFSB200::b200_push(b200md_ctx *ctx) { return B200_EAM_PUSH(); }
void PairEAMFSB200::compute(int eflag, int vflag) { B200_EAM_COMPUTE("eam/fs/b200"); }
