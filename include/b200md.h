/* b200md.h — C ABI of libb200md.so, the B200-native (sm_100a) short-range MD hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no C++/torch types.  A LAMMPS
 * suffix style (`lj/cut/b200`, `eam/b200`, `eam/alloy/b200`, `nve/b200`, `verlet/b200`; see
 * INTEGRATION.md) binds these exactly as the reference's GPU package binds its
 * `ljl_gpu_init / ljl_gpu_compute_n / ljl_gpu_clear` layer
 * (/root/reference/src/GPU/pair_lj_cut_gpu.cpp:36-54).
 *
 * Conventions
 *   - every call returns 0 on success, non-zero on failure; b200md_last_error() holds the
 *     message (the reference's Error::all/one text where one exists, src/error.cpp).
 *   - host pointers are borrowed for the duration of the call only; the context owns all
 *     device memory and its stream.
 *   - per-type tables are 1-based and contiguous, `(ntypes+1)*(ntypes+1)` doubles, like the
 *     reference's `memory->create(cutsq,n+1,n+1)` blocks (src/memory.h:144-160).
 *   - atom arrays on the host side are the reference's AoS `double x[n][3]`, `int type[n]`,
 *     `tagint tag[n]` (32-bit, LAMMPS_SMALLBIG, src/lmptype.h:90-92), `imageint image[n]`.
 *   - there is no CPU fallback: every compute entry point fails with
 *     "b200md: no CUDA device" when no sm_100 device is usable.
 */
#ifndef B200MD_H
#define B200MD_H

#include <stddef.h>

#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct b200md_ctx b200md_ctx;

#define B200MD_OK 0
#define B200MD_ERR 1

/* neighbor list kinds (reference: NeighRequest half/full, src/neigh_request.h:21-135) */
#define B200MD_LIST_FULL 0        /* NPairFullBinAtomonly, src/npair_full_bin_atomonly.cpp:32-92 */
#define B200MD_LIST_HALF_NEWTON 1 /* NPairHalfBinAtomonlyNewton, src/npair_half_bin_atomonly_newton.cpp:33-113 */

/* pair styles */
#define B200MD_PAIR_NONE 0
#define B200MD_PAIR_LJ_CUT 1 /* src/pair_lj_cut.cpp */
#define B200MD_PAIR_EAM 2    /* src/MANYBODY/pair_eam.cpp, pair_eam_alloy.cpp (same compute) */

const char *b200md_last_error(void);
const char *b200md_version(void);

/* ---- host-only helpers (no GPU needed) ------------------------------------------------ */

/* ProcMap::onelevel_grid + best_factors (src/procmap.cpp:46-132, 853-895): factor nranks
 * into px*py*pz minimising sub-domain surface for box lengths prd[3]. */
int b200md_procgrid(int nranks, const double prd[3], int procgrid[3]);
/* rank -> grid location, row-major like MPI_Cart_create without reorder (x slowest). */
int b200md_rank2loc(int rank, const int procgrid[3], int myloc[3]);
/* Domain::set_local_box (src/domain.cpp:335-357), uniform splits. */
int b200md_local_box(const double boxlo[3], const double boxhi[3], const int procgrid[3],
                     const int myloc[3], double sublo[3], double subhi[3]);

/* CommBrick::setup swap schedule (src/comm_brick.cpp:191-483), SINGLE mode, orthogonal box.
 * Arrays have room for B200MD_MAXSWAP entries. */
#define B200MD_MAXSWAP 64
typedef struct b200md_swap_schedule {
  int nswap;
  int maxneed[3];
  int dim[B200MD_MAXSWAP];       /* 0,1,2 */
  int ineed[B200MD_MAXSWAP];     /* index within the dim (even: send to -dim side) */
  int sendflag[B200MD_MAXSWAP];  /* ineed/2 < sendneed[dim][ineed%2] */
  int sendproc[B200MD_MAXSWAP];
  int recvproc[B200MD_MAXSWAP];
  int pbc[B200MD_MAXSWAP][3];    /* periodic shift multiples applied by pack_border/pack_comm */
  double slablo[B200MD_MAXSWAP];
  double slabhi[B200MD_MAXSWAP];
} b200md_swap_schedule;
int b200md_comm_schedule(const double boxlo[3], const double boxhi[3], const int periodic[3],
                         const int procgrid[3], const int myloc[3], double cutghost,
                         b200md_swap_schedule *out);

/* Direct halo tables (csrc/halo26.cu): the staged ghost set of CommBrick::borders (src/comm_brick.cpp:771-979) is the
 * product set over the dimensions of the per-dimension send slabs, so an owned atom can be classified once by the up
 * to 26 directions (dx,dy,dz), index (dx+1)+3(dy+1)+9(dz+1), it is a ghost in.  Per direction: the rank a copy lands
 * on (-1 = no neighbour), the periodic shift multiples (CommBrick::setup pbc flags); per dimension the slab bounds.
 * usable = 0 when some dimension needs more than one ghost layer (then the staged path runs). */
int b200md_halo_directions(const double boxlo[3], const double boxhi[3], const int periodic[3],
                           const int procgrid[3], const int myloc[3], double cutghost, int peer[27],
                           int shift[27][3], double slabhi_left[3], double slablo_right[3], int *usable);

/* NBinStandard::setup_bins (src/nbin_standard.cpp:82-214) + NStencil::create_setup
 * (src/nstencil.cpp:213-247): the reference's bin geometry for a sub-box. */
typedef struct b200md_bin_geometry {
  int nbinx, nbiny, nbinz;       /* global bins */
  int mbinx, mbiny, mbinz;       /* local bins incl. ghost layers */
  int mbinxlo, mbinylo, mbinzlo;
  int sx, sy, sz;                /* stencil half-extent */
  double binsizex, binsizey, binsizez;
  double bininvx, bininvy, bininvz;
} b200md_bin_geometry;
int b200md_bin_setup(const double boxlo[3], const double boxhi[3], const double sublo[3],
                     const double subhi[3], double cutghost, double cutneighmax,
                     b200md_bin_geometry *out);

/* fcc lattice + `velocity all create T seed loop geom` (src/create_atoms.cpp:1301-1370,
 * src/lattice.cpp:469-486, src/velocity.cpp:160-372, src/random_park.cpp:86-124) restricted
 * to what bench/in.lj and bench/in.eam use: `lattice fcc`, `region box block 0 nx 0 ny 0 nz`,
 * one type, dist uniform, mom yes, rot no.  `scale` is the lattice constant in box units
 * (lj: (4/rho*)^(1/3); metal: a).  Generates only the atoms inside [sublo,subhi).
 * Pass x=NULL to count.  Velocities are the un-scaled, un-zeroed draws; use
 * b200md_velocity_finalize after set_atoms (momentum zeroing and rescaling are global). */
int b200md_lattice_fcc(double scale, int nx, int ny, int nz, const double sublo[3],
                       const double subhi[3], int seed, double mass, int *natoms_out,
                       double *x_aos, double *v_aos, int *tag);

/* ---- context ------------------------------------------------------------------------- */

/* number of visible CUDA devices (hosts map rank -> device as rank % count, like
 * src/KOKKOS/kokkos.cpp:144-199) */
int b200md_device_count(int *count);
int b200md_create(int device, int rank, int nranks, b200md_ctx **out);
int b200md_destroy(b200md_ctx *ctx);

/* 128-byte NCCL unique id: rank 0 obtains it, the launcher broadcasts it (torch.distributed /
 * MPI), every rank then calls b200md_comm_init.  nranks==1 needs neither. */
int b200md_comm_unique_id(void *id128);
int b200md_comm_init(b200md_ctx *ctx, const void *id128);

/* unit constants (src/update.cpp:145-200): boltz, mvv2e, ftm2v, nktv2p */
int b200md_set_units(b200md_ctx *ctx, double boltz, double mvv2e, double ftm2v, double nktv2p);

/* Domain + Comm::set_procs: global box, periodicity, process grid (b200md_procgrid).  Computes
 * myloc / sub-box like the reference. */
int b200md_set_box(b200md_ctx *ctx, const double boxlo[3], const double boxhi[3],
                   const int periodic[3], const int procgrid[3]);
int b200md_get_subbox(b200md_ctx *ctx, double sublo[3], double subhi[3]);

/* Upload owned atoms (AtomVecAtomic fields, src/atom_vec_atomic.cpp).  mass is 1-based
 * [ntypes+1].  image may be NULL (all zero = (512,512,512) packed like Domain). */
int b200md_set_atoms(b200md_ctx *ctx, int nlocal, const double *x_aos, const double *v_aos,
                     const int *type, const int *tag, const int *image, int ntypes,
                     const double *mass);
/* Download owned atoms in current local order.  Any pointer may be NULL. */
int b200md_get_nlocal(b200md_ctx *ctx, int *nlocal, int *nghost);
int b200md_get_atoms(b200md_ctx *ctx, double *x_aos, double *v_aos, double *f_aos, int *type,
                     int *tag, int *image);
/* Overwrite x and/or v of the owned atoms IN CURRENT LOCAL ORDER (the order b200md_get_atoms
 * returns) from host arrays: the per-step upload of a host-resident integrator, like the
 * reference GPU package's per-step `host_x` (src/GPU/pair_lj_cut_gpu.cpp:98-102).  Asynchronous
 * on the context's stream when the host memory is pinned. */
int b200md_put_atoms(b200md_ctx *ctx, const double *x_aos, const double *v_aos);
/* Asynchronous download of x and/or f (owned atoms, current local order) into PINNED host arrays: packed on the
 * context's stream, copied on a separate copy stream, x first.  b200md_wait_download(ctx, 0) returns when x has
 * arrived, (ctx, 1) when everything has.  A host that re-uploads positions every step (b200md_put_atoms) can do so as
 * soon as x is there; the f copy then shares the PCIe link's other direction with that upload and overlaps the next
 * step's kernels — the per-step boundary of the reference's GPU package (host x up, f down,
 * src/GPU/pair_lj_cut_gpu.cpp:98-130) with the copies double-buffered. */
int b200md_get_atoms_async(b200md_ctx *ctx, double *x_aos, double *f_aos);
int b200md_wait_download(b200md_ctx *ctx, int all);
/* The same boundary, streamed: b200md_get_atoms_async sends x in up to B200MD_DL_PARTS pieces of whole atoms.
 * b200md_wait_download_part(ctx, p, &first, &count) returns when piece p (atoms first .. first+count-1) is on the
 * host (count = 0 beyond the last piece); b200md_put_atoms_range uploads the positions of that range from the host
 * array x_aos (row 0 = atom 0) on the context's stream.  Turning piece p around while piece p+1 is still coming
 * down uses both directions of the link at once (the reference's GPU package overlaps its copies with kernels the
 * same way, lib/gpu/lal_atom.h add_x_data / lal_answer.cpp copy_answers on their own command queues). */
#define B200MD_DL_PARTS 8
int b200md_wait_download_part(b200md_ctx *ctx, int part, int *first, int *count);
int b200md_put_atoms_range(b200md_ctx *ctx, const double *x_aos, int first, int count);
/* pinned host memory for put/get */
int b200md_host_alloc(void **ptr, size_t bytes);
int b200md_host_free(void *ptr);
/* ghost atoms (x with periodic shift applied, tag, type) in local order nlocal..nall-1 */
int b200md_get_ghosts(b200md_ctx *ctx, double *x_aos, int *type, int *tag);

/* Velocity::zero_momentum + Velocity::rescale (src/velocity.cpp:358-368, 663-700 region):
 * global over all ranks. */
int b200md_velocity_finalize(b200md_ctx *ctx, double t_desired, int zero_momentum);

/* ---- neighbor ------------------------------------------------------------------------ */

/* Neighbor::init cutoffs (src/neighbor.cpp:348-388) + neigh_modify every/delay/check
 * (src/neighbor.cpp:2622-…).  cutneighsq is (ntypes+1)^2.  list_kind = B200MD_LIST_*. */
int b200md_neigh_config(b200md_ctx *ctx, int ntypes, const double *cutneighsq,
                        double cutneighmax, double skin, int every, int delay, int check,
                        int list_kind);
/* Neighbor::decide (src/neighbor.cpp:2309-2324) incl. check_distance (:2338-2390). */
int b200md_neigh_decide(b200md_ctx *ctx, int *rebuild);
/* The reneighbor branch of Verlet::run (src/verlet.cpp:289-307): Domain::pbc, Comm::exchange,
 * spatial sort, Comm::borders, Neighbor::build (bin_atoms + NPair::build). */
int b200md_neigh_build(b200md_ctx *ctx);
/* non-reneighbor branch: Comm::forward_comm (src/comm_brick.cpp:538-591). */
int b200md_forward_comm(b200md_ctx *ctx);
/* Export the list for parity checks.  numneigh[nlocal]; neigh_flat receives, atom after
 * atom, the neighbors as local indices (by_tag=0) or as tags (by_tag=1), unsorted.
 * Pass neigh_flat=NULL to obtain only numneigh and *total. */
int b200md_neigh_export(b200md_ctx *ctx, int by_tag, int *numneigh, int *neigh_flat,
                        long long *total);
/* Storage layout of the current list: *cluster = 1 when 4 consecutive owned atoms share one row of j entries with
 * 4-bit pair masks (atomic systems, option cluster), 0 for per-atom rows; *entries = stored entries over all rows
 * (what the force kernel gathers per step); *lanes = lanes that walk one row.  Any pointer may be NULL. */
int b200md_neigh_layout(b200md_ctx *ctx, int *cluster, long long *entries, int *lanes);
int b200md_neigh_stats(b200md_ctx *ctx, long long *nbuilds, long long *ndanger,
                       long long *total_neighbors, int *maxneigh_pitch);

/* ---- pair ---------------------------------------------------------------------------- */

/* PairLJCut::init_one products (src/pair_lj_cut.cpp:501-556): all (ntypes+1)^2, 1-based. */
int b200md_pair_lj_cut_config(b200md_ctx *ctx, int ntypes, const double *cutsq,
                              const double *lj1, const double *lj2, const double *lj3,
                              const double *lj4, const double *offset);
/* PairEAM::array2spline products (src/MANYBODY/pair_eam.cpp:744-797).  Splines are the
 * reference's `[ntable][n+1][7]` blocks (row 0 unused, 1-based rows); type2* are 1-based
 * `[ntypes+1]` / `[(ntypes+1)^2]` maps. */
int b200md_pair_eam_config(b200md_ctx *ctx, int ntypes, int nrho, int nr, double rdr,
                           double rdrho, double rhomax, double cutforcesq, int nfrho,
                           const double *frho_spline, int nrhor, const double *rhor_spline,
                           int nz2r, const double *z2r_spline, const int *type2frho,
                           const int *type2rhor, const int *type2z2r, const double *scale);
/* Pair::compute(eflag,vflag) (src/pair_lj_cut.cpp:69-137, src/MANYBODY/pair_eam.cpp:142-328);
 * includes Verlet::force_clear (fused) and, for EAM, the fp forward_comm
 * (pair_eam.cpp:251).  eng_vdwl / virial[6] (xx,yy,zz,xy,xz,yz) are this rank's
 * contributions, written only when the corresponding flag is set. */
int b200md_pair_compute(b200md_ctx *ctx, int eflag, int vflag, double *eng_vdwl,
                        double virial[6]);
/* This rank's Pair::eng_vdwl and Pair::virial[6] (src/pair.h:45-46) as left by the last force
 * evaluation that tallied them (b200md_setup, an output step of b200md_run, or
 * b200md_pair_compute with flags set). */
int b200md_pair_tallies(b200md_ctx *ctx, double *eng_vdwl, double virial[6]);
/* Comm::reverse_comm of forces (src/comm_brick.cpp:598-639); no-op for full lists. */
int b200md_reverse_comm(b200md_ctx *ctx);

/* ---- molecular topology (bench/in.chain) ------------------------------------------------ */

/* Bonds and special neighbors of an `atom_style bond` system, keyed by atom TAG and replicated on
 * every rank (nothing per-atom migrates).  CSR tables: *_offset has natoms+2 entries, row t (1-based
 * tag) spans [offset[t], offset[t+1]).  bond_* lists the bonds STORED on tag t (newton_bond on:
 * each bond once, src/atom_vec_bond.cpp data_atom / Atom::data_bonds src/atom.cpp:1408-1470);
 * special_* is Atom::special with its level (1 = 1-2, 2 = 1-3, 3 = 1-4; src/special.cpp:50-120).
 * special_lj / special_coul are Force::special_lj/coul[4] (index 0 unused); weights must be 0 or
 * 1 (special_bonds fene/lj 0 1 1: 1-2 pairs leave the pair list, src/npair.h:125-160).  The bond
 * list of the reference (NTopoBondAll::build, src/ntopo_bond_all.cpp:37-80) is implicit: each
 * owned atom evaluates the bonds stored on its tag through a tag->index map rebuilt with the
 * neighbor list. */
int b200md_set_topology(b200md_ctx *ctx, long long natoms, const int *bond_offset,
                        const int *bond_partner_tag, const int *bond_type,
                        const int *special_offset, const int *special_tag,
                        const int *special_level, const double special_lj[4],
                        const double special_coul[4]);
/* bond_style fene coefficients, 1-based [nbondtypes+1] (BondFENE::coeff,
 * src/MOLECULE/bond_fene.cpp:146-170). */
int b200md_bond_fene_config(b200md_ctx *ctx, int nbondtypes, const double *k, const double *r0,
                            const double *epsilon, const double *sigma);
/* Bond::compute(eflag,vflag) (src/MOLECULE/bond_fene.cpp:53-129): adds the bond forces to f
 * (call after b200md_pair_compute, before b200md_reverse_comm); energy / virial[6] are this
 * rank's Bond::energy and Bond::virial, written when the flag is set. */
int b200md_bond_compute(b200md_ctx *ctx, int eflag, int vflag, double *energy, double virial[6]);
/* This rank's Bond::energy and Bond::virial[6] (src/bond.h:32-33) of the last tallied evaluation. */
int b200md_bond_tallies(b200md_ctx *ctx, double *energy, double virial[6]);

/* ---- per-atom energy / virial (output steps) -------------------------------------------- */

#define B200MD_PERATOM_PAIR 1 /* force->pair->eatom / vatom */
#define B200MD_PERATOM_BOND 2 /* force->bond->eatom / vatom */
/* The eflag_atom / vflag_atom tallies of Pair::ev_tally (src/pair.cpp:1036-1040, 1078-1097),
 * PairEAM's eatom[i] += phi (src/MANYBODY/pair_eam.cpp:245) and Bond::ev_tally
 * (src/bond.cpp:186-192, 234-251), summed over the styles selected by `which` and with ghost
 * contributions already returned to their owners — i.e. what ComputePEAtom::compute_peratom
 * (src/compute_pe_atom.cpp:92-170) leaves in `energy[]` and ComputeStressAtom::compute_peratom
 * (src/compute_stress_atom.cpp:140-240) in `stress[][]` BEFORE its kinetic term and its
 * -nktv2p factor.  Evaluated on the device for the current positions and neighbor list; the
 * forces must be current (after b200md_setup, b200md_run or b200md_pair_compute at these
 * positions: EAM reuses their rho / fp).  Forces and global tallies are not touched.
 * eatom: host [nlocal] or NULL; vatom: host [nlocal][6] (xx yy zz xy xz yz) or NULL; rows in
 * the device's local order, the order b200md_get_atoms returns. */
int b200md_peratom(b200md_ctx *ctx, int eflag_atom, int vflag_atom, int which, double *eatom, double *vatom);

/* ---- fix langevin ---------------------------------------------------------------------- */

/* `fix ID all langevin Tstart Tstop damp seed [scale type ratio ...]` (src/fix_langevin.cpp:60-150,
 * defaults tally no / zero no / gjf no).  After every force evaluation — the set-up one included,
 * FixLangevin::setup :360-361 — each owned atom gets f += gamma1 v + gamma2 (u - 0.5) with
 * gamma1 = -m/damp/ftm2v/ratio, gamma2 = sqrt(m) sqrt(24 boltz/damp/dt/mvv2e)/ftm2v/sqrt(ratio)
 * sqrt(T_target) (init :286-298, post_force_templated :636-713) and T_target ramped from Tstart
 * to Tstop over the run (compute_target :773-787; the window is the nsteps of each b200md_run
 * unless b200md_fix_langevin_window is used).  u are RanMars::uniform() draws
 * (src/random_mars.cpp:78-95), three per atom in the HOST's local atom order: the order of the
 * last b200md_set_atoms call, which a caller that wants the reference's numbers makes in
 * Atom::sort order (b200md_atom_sort_order).  The device reproduces the serial stream exactly:
 * each step's 3*nlocal draws are generated in parallel chunks by jump-ahead of the generator's
 * linear recurrence.  One rank only (the reference seeds rank r with seed + r and draws in that
 * rank's local order, which no other decomposition can replay); refused otherwise.
 * ratio: [ntypes+1] or NULL (all 1); dt: the timestep of the set-up evaluation (update->dt at
 * FixLangevin::init; b200md_run takes over its own dt like FixLangevin::reset_dt).
 * Call after b200md_set_atoms and before b200md_setup. */
int b200md_fix_langevin(b200md_ctx *ctx, double t_start, double t_stop, double t_period, int seed,
                        const double *ratio, double dt);
/* Update::beginstep / endstep of the enclosing `run` and the current timestep, for callers that
 * split one run into several b200md_run calls (verlet/b200 stops at output steps).  endstep ==
 * beginstep switches back to "every b200md_run is its own window". */
int b200md_fix_langevin_window(b200md_ctx *ctx, long long ntimestep, long long beginstep, long long endstep);
/* Atom::sortfreq of the host (`atom_modify sort`, default 1000; 0 = never).  The reference re-sorts its
 * atoms at the first reneighboring at or after every multiple of it (Verlet::run, src/verlet.cpp:285;
 * Atom::sort sets nextsort, src/atom.cpp:2252), which changes the order the draws are handed out in; the
 * device re-ranks its stream index the same way at that rebuild (stable by bin, previous rank inside a
 * bin).  The host's own arrays are not re-ordered by this. */
int b200md_fix_langevin_sort(b200md_ctx *ctx, int sortfreq);
/* `unfix`: stop applying it. */
int b200md_fix_langevin_off(b200md_ctx *ctx);

/* ---- RanMars stream, Atom::sort order (host-side helpers, no GPU needed) ------------------ */

#define B200MD_RANMARS_LAG 97
/* State of RanMars(seed) right after construction (src/random_mars.cpp:30-72), in integer form:
 * the 97 most recent lagged-Fibonacci values in units of 2^-24, oldest first. */
int b200md_ranmars_state(int seed, uint32_t state[B200MD_RANMARS_LAG]);
/* The state k draws later (x^k mod x^97 + x^64 - 1 over Z/2^24 applied to `state`). */
int b200md_ranmars_jump(const uint32_t state[B200MD_RANMARS_LAG], long long k, uint32_t out[B200MD_RANMARS_LAG]);
/* `count` consecutive RanMars::uniform() values from `state`; position = number of draws made
 * before it (1 right after construction), needed for the generator's arithmetic sequence c. */
int b200md_ranmars_draws(const uint32_t state[B200MD_RANMARS_LAG], long long position, int count, double *out);
/* Device generator check: `fills` consecutive fills of n draws each, exactly as fix langevin
 * requests them per force evaluation; out: host [fills*n]. */
int b200md_ranmars_device_stream(b200md_ctx *ctx, int seed, int n, int fills, double *out);
/* Atom::sort's permutation (src/atom.cpp:2246-2349, bins of cutneighmax/2 over the sub-box,
 * stable): new atom I = old atom permute[I].  x: [n][3] wrapped positions. */
int b200md_atom_sort_order(int n, const double *x, const double sublo[3], const double subhi[3],
                           double cutneighmax, int *permute);

/* ---- fix nve + thermo ---------------------------------------------------------------- */

/* FixNVE::initial_integrate / final_integrate (src/fix_nve.cpp:64-141). dtf = 0.5*dt*ftm2v. */
int b200md_nve_initial(b200md_ctx *ctx, double dtv, double dtf);
int b200md_nve_final(b200md_ctx *ctx, double dtf);
/* ComputeTemp / ComputePE / ComputePressure scalars (src/compute_temp.cpp:72-99,
 * src/compute_pe.cpp:84-112, src/compute_pressure.cpp:228-262), summed over ranks.
 * Requires that the last pair_compute ran with eflag=vflag=1. */
typedef struct b200md_thermo {
  long long step;
  double natoms;
  double temp, pe, ke, press; /* pe, ke extensive (not divided by N); pe = pair + bond */
  double virial[6];           /* pair + bond */
  double emol;                /* bond energy (thermo keyword emol), 0 for atomic systems */
} b200md_thermo;
int b200md_thermo_compute(b200md_ctx *ctx, b200md_thermo *out);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under stream capture, tight serving loops, harnesses that forbid it).  The kinetic sums behind
 * b200md_thermo_compute's temp / ke — out5 = {sum m v.v, sum m vx, sum m vy, sum m vz, sum m} over n atoms — in ONE
 * launch that allocates nothing: the per-block partial sums go to a static device array inside the library
 * (164 096 bytes, held by every GPU that loads the library until exit) instead of the context's growable scratch.
 * All pointers are DEVICE pointers of the CURRENT device, owned by the caller: v_soa_dev = velocities as
 * [3][stride] (component k of atom i at k*stride + i, stride >= n), type_dev = atom types 1..ntypes (NULL: all
 * type 1), mass_dev = mass by type [ntypes+1] (Atom::mass), out5_dev = 5 doubles.  The launch goes to `stream`
 * (cudaStream_t as void*, NULL = default stream) and is not synchronised.  n = 0 only does the once-per-device
 * set-up (the static array's device address): call that once outside a stream capture.
 * Limits: n <= B200MD_KE_STATIC_MAX_ATOMS, else B200MD_ERR with a message naming the limit.  The array is shared
 * by all calls on a device: calls that overlap on the GPU (streams with no ordering between them) race; calls
 * ordered on one stream, or by events, are safe.  Same input, same n -> same bits on every call. */
#define B200MD_KE_STATIC_MAX_ATOMS 16777216
int b200md_ke_sums_static(int n, long long stride, const double *v_soa_dev, const int *type_dev,
                          const double *mass_dev, double *out5_dev, void *stream);

/* ---- whole loop ---------------------------------------------------------------------- */

/* Verlet::setup (src/verlet.cpp:93-162): initial build + force evaluation with energy/virial. */
int b200md_setup(b200md_ctx *ctx);
/* Verlet::run(nsteps) (src/verlet.cpp:229-360) with fix nve, thermo every `thermo_every`
 * steps (0: only last step).  thermo_out (may be NULL) receives up to max_thermo records
 * (the step-0 record is produced by b200md_setup + b200md_thermo_compute). */
int b200md_run(b200md_ctx *ctx, int nsteps, double dt, int thermo_every,
               b200md_thermo *thermo_out, int max_thermo, int *n_thermo);

/* timing / counters of the last b200md_run (device time by CUDA events on the context's
 * stream, ms): total, pair, neigh(build incl. sort/borders), comm, modify(nve); kernel
 * launches issued. */
typedef struct b200md_timing {
  double total_ms, pair_ms, neigh_ms, comm_ms, modify_ms;
  long long launches;
  long long pair_launches;
  int nbuilds;
} b200md_timing;
int b200md_get_timing(b200md_ctx *ctx, b200md_timing *out);
/* tuning knob: lanes cooperating on one atom in the pair kernels (1,2,4,8,16,32); the list
 * build has its own lane count (default 1). Call before b200md_setup. */
int b200md_set_lanes_per_atom(b200md_ctx *ctx, int lanes);
/* Runtime options, the role the reference gives its `package <accelerator> <keyword> <value>`
 * command (Input::package src/input.cpp -> KokkosLMP::accelerator src/KOKKOS/kokkos.cpp:300-560,
 * FixGPU src/GPU/fix_gpu.cpp:90-230): `package b200 <keyword> <value>` in verlet/b200.  The library
 * reads NO environment variables.  Keywords (value as text):
 *   halo         direct | staged   per-step ghost traffic owner<->holder in one exchange, or the
 *                                  reference's three forwarding stages (several ranks only)
 *   overlap      on | off          compute / halo overlap on several ranks
 *   split        0..1              share of the interior tiles evaluated before the boundary tiles
 *   fuse_nve     on | off          final_integrate(n) + initial_integrate(n+1) in one pass (default on)
 *   eam_smem     on | off | auto   one-element EAM tables staged in shared memory, rows swizzled against bank conflicts
 *                                  (auto, the default: the force pass only; on: every pass)
 *   eam_mask     on | off          half list: the density pass marks the list entries inside the force cutoff, the force
 *                                  pass walks only those (default on)
 *   bin_scale    auto | 0.25..4    neighbor bins of bin_scale x cutneighmax/2 (1 = NBinStandard's size; list membership is
 *                                  decided by distance, so this is a device-side choice).  auto: 1.2 for cluster rows,
 *                                  0.6 for per-atom rows of atomic systems, 2 for molecular systems
 *   cell_order   x | index         order of the atoms inside a bin: ascending x (default) or atom index / tag
 *   self_border  direct | staged   one rank: all ghost images of a rebuild in one pass (26 directions), or the
 *                                  reference's three stages (default direct; staged is used when a side needs
 *                                  more than one ghost layer)
 *   phase_timing on | off          per-phase CUDA events in b200md_run (b200md_get_timing)
 *   trace        on | off          timestamps of both streams' milestones of an overlapped step on stderr (development)
 *   lanes        1,2,...,32        = b200md_set_lanes_per_atom
 *   build_lanes  1,2,...,32        lanes per atom in the list build
 *   cluster      on | off | auto   cluster-pair neighbor list for atomic systems: 4 consecutive atoms of the bin-sorted
 *                                  order share one j-list whose entries carry a 4-bit pair mask (auto, the default:
 *                                  lj/cut yes, eam no)
 *   graph        on | off          small-system mode: non-rebuild steps replayed as a CUDA graph
 * Unknown keywords and bad values fail with "Illegal package b200 command: ...". */
int b200md_set_option(b200md_ctx *ctx, const char *key, const char *value);
int b200md_get_option(b200md_ctx *ctx, const char *key, char *value, int len);
/* cudaStream_t of the context (as void*), for callers that time with their own events. */
int b200md_get_stream(b200md_ctx *ctx, void **stream);

#ifdef __cplusplus
}
#endif
#endif
