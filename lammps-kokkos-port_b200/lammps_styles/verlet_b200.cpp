// run_style verlet/b200 — see verlet_b200.h.  Written against the reference's public class
// interfaces (Integrate/Verlet, Atom, Neighbor, Force, Modify, Output); the physics is in
// libb200md.so.

#include "verlet_b200.h"

#include "atom.h"
#include "atom_vec.h"
#include "bond.h"
#include "comm.h"
#include "domain.h"
#include "error.h"
#include "fix.h"
#include "force.h"
#include "modify.h"
#include "neigh_list.h"
#include "neigh_request.h"
#include "neighbor.h"
#include "npair.h"
#include "npair_b200.h"
#include "output.h"
#include "pair.h"
#include "timer.h"
#include "update.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>

using namespace LAMMPS_NS;

b200md_ctx *B200Shared::ctx = nullptr;

/* ---------------------------------------------------------------------- */

VerletB200::VerletB200(LAMMPS *lmp, int narg, char **arg) :
    Verlet(lmp, narg, arg), ctx(nullptr), host_stale(false), builds_at_setup(0), danger_at_setup(0), ghosts_step(-1)
{
  // run_style verlet/b200 [keyword value ...]: the runtime options of the device library, the role
  // `package gpu|kokkos <keyword> <value>` plays for the reference's accelerator packages (an unmodified
  // Input::package knows only its own package names, src/input.cpp).  Keywords: include/b200md.h,
  // b200md_set_option (halo, overlap, split, fuse_nve, eam_smem, lanes, build_lanes, graph).
  if (narg % 2) error->all(FLERR, "Illegal run_style verlet/b200 command: keyword without a value");
  for (int i = 0; i + 1 < narg; i += 2) options.emplace_back(arg[i], arg[i + 1]);
}

VerletB200::~VerletB200()
{
  if (ctx) {
    b200md_destroy(ctx);
    if (B200Shared::ctx == ctx) B200Shared::ctx = nullptr;
  }
}

void VerletB200::fail()
{
  error->all(FLERR, b200md_last_error());
}

/* ----------------------------------------------------------------------
   what the device loop covers: SURVEY.md section 8 -- atomic systems, orthogonal periodic
   or non-periodic 3-d box, one pair style with a b200 variant, fix nve on all atoms
------------------------------------------------------------------------- */

void VerletB200::check_supported()
{
  if (domain->triclinic) error->all(FLERR, "Run style verlet/b200 requires an orthogonal box");
  if (domain->dimension != 3) error->all(FLERR, "Run style verlet/b200 requires a 3d system");
  if (force->angle || force->dihedral || force->improper || force->kspace || atom->molecular == Atom::TEMPLATE)
    error->all(FLERR, "Run style verlet/b200 supports pair styles and bond styles only");
  if (atom->molecular != Atom::ATOMIC) {
    // bench/in.chain: bonds only, newton_bond on, one rank (the host arrays hold the whole topology)
    if (force->bond && !dynamic_cast<B200BondStyle *>(force->bond))
      error->all(FLERR, "Run style verlet/b200 requires a bond style with the b200 suffix");
    if (!force->newton_bond) error->all(FLERR, "Run style verlet/b200 requires newton_bond on");
    if (comm->nprocs > 1) error->all(FLERR, "Run style verlet/b200: molecular systems run on one rank");
    if (!atom->tag_consecutive()) error->all(FLERR, "Run style verlet/b200 requires consecutive atom IDs");
  }
  if (atom->rmass_flag || atom->torque_flag)
    error->all(FLERR, "Run style verlet/b200 requires per-type masses and point particles");
  if (!force->pair || !dynamic_cast<B200PairStyle *>(force->pair))
    error->all(FLERR, "Run style verlet/b200 requires a pair style with the b200 suffix");
  if (domain->box_change) error->all(FLERR, "Run style verlet/b200 requires a fixed box");

  int nintegrate = 0;
  for (const auto &fix : modify->get_fix_list()) {
    if (fix->time_integrate) {
      if (!utils::strmatch(fix->style, "^nve/b200$") || fix->igroup != 0)
        error->all(FLERR, "Run style verlet/b200 requires fix nve/b200 on group all");
      nintegrate++;
    }
  }
  // `run 0` (forces and tallies only) needs no integrator
  if (nintegrate > 1 || (nintegrate == 0 && update->nsteps > 0))
    error->all(FLERR, "Run style verlet/b200 requires exactly one fix nve/b200");
  // fixes that act after the force evaluation: only those the device loop applies itself
  int ndevice_post = 0;
  for (const auto &fix : modify->get_fix_list())
    if (dynamic_cast<B200FixStyle *>(fix)) ndevice_post++;
  if (ndevice_post > 1) error->all(FLERR, "Run style verlet/b200 supports one fix langevin/b200");
  if (modify->n_initial_integrate != nintegrate || modify->n_final_integrate != nintegrate ||
      modify->n_post_integrate || modify->n_pre_exchange || modify->n_pre_neighbor ||
      modify->n_post_neighbor || modify->n_pre_force || modify->n_pre_reverse ||
      modify->n_post_force_any != ndevice_post)
    error->all(FLERR, "Run style verlet/b200 does not support fixes that act during the step");
  // end_of_step fixes that only READ the state (fix ave/time, ave/atom, ave/chunk, ave/histo, ave/correlate, vector,
  // print) are served: the device loop stops on the steps they are due, the host arrays, tallies and lists are
  // refreshed, and Modify::end_of_step runs like in Verlet::run (verlet.cpp:349).  One that writes atom data on the
  // host (fix momentum, ...) would be lost on the device: refused.
  for (int i = 0; i < modify->nfix; i++)
    if ((modify->fmask[i] & FixConst::END_OF_STEP) &&
        !utils::strmatch(modify->fix[i]->style, "^ave/") && !utils::strmatch(modify->fix[i]->style, "^vector$") &&
        !utils::strmatch(modify->fix[i]->style, "^print$"))
      error->all(FLERR, "Run style verlet/b200 does not support fix {} (it acts on the host's atoms during the step)",
                 modify->fix[i]->style);
}

/* ---------------------------------------------------------------------- */

void VerletB200::init()
{
  Verlet::init();

  if (!ctx) {
    int ngpu = 0;
    b200md_device_count(&ngpu);
    if (ngpu < 1) error->all(FLERR, "Run style verlet/b200: no CUDA device");
    if (b200md_create(comm->me % ngpu, comm->me, comm->nprocs, &ctx)) fail();
    if (comm->nprocs > 1) {
      char id[128];
      memset(id, 0, sizeof(id));
      if (comm->me == 0 && b200md_comm_unique_id(id)) fail();
      MPI_Bcast(id, 128, MPI_BYTE, 0, world);
      if (b200md_comm_init(ctx, id)) fail();
    }
    B200Shared::ctx = ctx;
    for (const auto &kv : options)
      if (b200md_set_option(ctx, kv.first.c_str(), kv.second.c_str())) fail();
  }
  if (b200md_set_units(ctx, force->boltz, force->mvv2e, force->ftm2v, force->nktv2p)) fail();
  if (b200md_set_box(ctx, domain->boxlo, domain->boxhi, domain->periodicity, comm->procgrid)) fail();
}

/* ----------------------------------------------------------------------
   upload atoms + settings, first build, first force evaluation with tallies
------------------------------------------------------------------------- */

void VerletB200::device_setup()
{
  check_supported();    // Modify::init has filled its per-hook fix lists by now
  atom->setup();
  domain->pbc();
  domain->reset_box();
  comm->setup();
  // the local order Verlet::setup leaves (src/verlet.cpp:123): it decides which RanMars draws an
  // atom gets from fix langevin; nothing else on the device depends on it
  if (atom->sortfreq > 0 && comm->nprocs == 1) {
    atom->nghost = 0;
    atom->sort();
    if (atom->map_style != Atom::MAP_NONE) {
      atom->map_init();
      atom->map_set();
    }
  }

  const int nlocal = atom->nlocal;
  const int *mask = atom->mask;
  bool uniform = true;
  for (int i = 1; i < nlocal; i++)
    if (mask[i] != mask[0]) { uniform = false; break; }
  mask_by_tag.clear();
  if (!uniform) {
    mask_by_tag.reserve(nlocal);
    for (int i = 0; i < nlocal; i++) mask_by_tag.emplace_back((int) atom->tag[i], mask[i]);
    std::sort(mask_by_tag.begin(), mask_by_tag.end());
  }

  host_index.clear();
  if (comm->nprocs == 1) {
    host_index.reserve(nlocal);
    for (int i = 0; i < nlocal; i++) host_index.emplace_back((int) atom->tag[i], i);
    std::sort(host_index.begin(), host_index.end());
  }

  const double *x = nlocal ? &atom->x[0][0] : nullptr;
  const double *v = nlocal ? &atom->v[0][0] : nullptr;
  if (b200md_set_atoms(ctx, nlocal, x, v, atom->type, atom->tag, atom->image, atom->ntypes, atom->mass))
    fail();
  // `newton on` (the default) -> half list with f_j / rho_j accumulated by FP64 RED and
  // reverse_comm (the reference's half/bin/atomonly/newton, chosen from the same flag in
  // Neighbor::choose_pair, src/neighbor.cpp:2073-2077); `newton off` -> full list, no atomics,
  // no reverse_comm (same forces as the reference's newtoff half list)
  if (b200md_neigh_config(ctx, atom->ntypes, &neighbor->cutneighsq[0][0], neighbor->cutneighmax,
                          neighbor->skin, neighbor->every, neighbor->delay, neighbor->dist_check,
                          force->newton_pair ? B200MD_LIST_HALF_NEWTON : B200MD_LIST_FULL))
    fail();
  if (dynamic_cast<B200PairStyle *>(force->pair)->b200_push(ctx)) fail();
  if (atom->molecular != Atom::ATOMIC) push_topology();
  bool thermostat = false;
  for (const auto &fix : modify->get_fix_list())
    if (auto dev = dynamic_cast<B200FixStyle *>(fix)) {
      if (dev->b200_push_fix(ctx)) fail();
      thermostat = true;
    }
  if (!thermostat && b200md_fix_langevin_off(ctx)) fail();    // `unfix` between runs
  if (b200md_setup(ctx)) fail();

  long long nb = 0, nd = 0;
  b200md_neigh_stats(ctx, &nb, &nd, nullptr, nullptr);
  builds_at_setup = nb;
  danger_at_setup = nd;
  neighbor->ncalls = 0;
  neighbor->ndanger = 0;

  force->setup();
  ev_set(update->ntimestep);
  host_stale = true;
  sync_host();
  publish_lists();
  publish_peratom();    // clears the styles' global accumulators too (ev_setup): before the tallies
  publish_tallies();
}

void VerletB200::setup(int flag)
{
  if (comm->me == 0 && screen) {
    fputs("Setting up Verlet/b200 run ...\n", screen);
    if (flag) {
      fmt::print(screen, "  Unit style    : {}\n  Current step  : {}\n  Time step     : {}\n",
                 update->unit_style, update->ntimestep, update->dt);
      timer->print_timeout(screen);
    }
  }
  update->setupflag = 1;
  device_setup();
  modify->setup(vflag);
  output->setup(flag);
  update->setupflag = 0;
}

void VerletB200::setup_minimal(int /*flag*/)
{
  update->setupflag = 1;
  device_setup();
  modify->setup(vflag);
  update->setupflag = 0;
}

/* ----------------------------------------------------------------------
   Pair::eng_vdwl / Pair::virial of this rank, read by compute pe / compute pressure
------------------------------------------------------------------------- */

void VerletB200::publish_tallies()
{
  Pair *pair = force->pair;
  pair->eng_vdwl = pair->eng_coul = 0.0;
  if (b200md_pair_tallies(ctx, &pair->eng_vdwl, pair->virial)) fail();
  if (force->bond && atom->molecular != Atom::ATOMIC) {
    if (b200md_bond_tallies(ctx, &force->bond->energy, force->bond->virial)) fail();
  }
}

/* ----------------------------------------------------------------------
   Pair::eatom / vatom and Bond::eatom / vatom on steps where a compute asked for them
   (Integrate::ev_set, src/integrate.cpp:131-183), read by compute pe/atom and stress/atom
   (src/compute_pe_atom.cpp:92-170, src/compute_stress_atom.cpp:140-340).  The device returns
   rows with the ghost contributions already summed into their owners; those computes still run
   their own reverse_comm when newton is on, so the host gets a consistent ghost section
   (Comm::borders: swap tables + ghost rows, which ev_setup leaves at zero).
------------------------------------------------------------------------- */

/* ----------------------------------------------------------------------
   host ghosts of the current step: Comm::borders on the synced positions, once per step
------------------------------------------------------------------------- */

void VerletB200::host_borders()
{
  if (ghosts_step == update->ntimestep && atom->nghost > 0) return;
  if (atom->map_style != Atom::MAP_NONE) atom->map_clear();
  atom->nghost = 0;
  comm->borders();
  ghosts_step = update->ntimestep;
}

/* ----------------------------------------------------------------------
   NeighList publication (NPair::build, src/neighbor.cpp:2470-2476).  Neighbor::build never runs under this run
   style, so a host-side consumer of a neighbor list (compute coord/atom, compute rdf, ...) would see nothing.  On
   output steps every plain bin list of the host — the pair style's own perpetual list included, occasional lists
   are often copies of it — is filled by NPairFullBinB200 / NPairHalfBinNewtonB200 from the device's list, and the
   reference's own NPair objects are marked up to date so Neighbor::build_one (src/neighbor.cpp:2528-2572) leaves
   the lists alone (NeighList::np is that object).  The request flags are protected members of the unmodified
   NeighRequest: read through pointers to members.
------------------------------------------------------------------------- */

namespace {
struct RequestAccess : public NeighRequest {
  static int NeighRequest::*p_half() { return &RequestAccess::half; }
  static int NeighRequest::*p_skip() { return &RequestAccess::skip; }
  static int NeighRequest::*p_halffull() { return &RequestAccess::halffull; }
};
}    // namespace

void VerletB200::publish_lists()
{
  bool wanted = false;
  for (int i = 0; i < neighbor->nlist; i++)
    if (neighbor->lists[i]->requestor != (void *) force->pair) wanted = true;
  if (!wanted) return;
  if (comm->nprocs > 1) error->all(FLERR, "Run style verlet/b200: host-side neighbor lists are published on one rank only");
  host_borders();
  static NPairFullBinB200 *npfull = nullptr;
  static NPairHalfBinNewtonB200 *nphalf = nullptr;
  static LAMMPS *owner = nullptr;
  if (owner != lmp) { npfull = new NPairFullBinB200(lmp); nphalf = new NPairHalfBinNewtonB200(lmp); owner = lmp; }
  const bool dbg = getenv("B200_DEBUG_LISTS") != nullptr;
  for (int pass = 0; pass < 2; pass++)
    for (int i = 0; i < neighbor->nlist; i++) {
      NeighList *list = neighbor->lists[i];
      NeighRequest *rq = neighbor->old_requests[i];    // Neighbor::init hands the requests of this run over to old_requests
      if ((pass == 1) != (list->copy != 0)) continue;    // parents first, then the lists that alias them
      if (dbg) { fprintf(stderr, "[lists] step %ld list %d copy %d occasional %d np %p\n", (long) update->ntimestep, i, list->copy, list->occasional, (void *) list->np); fflush(stderr); }
      if (list->copy) {                                   // NPairCopy (list->np): takes the parent's arrays
        if (list->np) list->np->build(list);
      } else {
        if (rq->*RequestAccess::p_skip() || rq->*RequestAccess::p_halffull() || list->ghost || list->respaouter)
          error->all(FLERR, "Run style verlet/b200 publishes plain half and full neighbor lists only");
        if (rq->*RequestAccess::p_half()) nphalf->build(list);
        else npfull->build(list);
      }
      if (list->np) list->np->last_build = update->ntimestep;
    }
  if (dbg) { fprintf(stderr, "[lists] done\n"); fflush(stderr); }
}

void VerletB200::publish_peratom()
{
  if (!(eflag & ENERGY_ATOM) && !(vflag & (VIRIAL_ATOM | VIRIAL_CENTROID))) return;
  if (vflag & VIRIAL_CENTROID)
    error->all(FLERR, "Run style verlet/b200 does not tally the centroid per-atom stress");
  const int nlocal = atom->nlocal;
  std::vector<int> rows(nlocal);
  if (comm->nprocs == 1 && (int) host_index.size() == nlocal && (int) htag.size() == nlocal) {
    for (int k = 0; k < nlocal; k++)
      rows[k] = std::lower_bound(host_index.begin(), host_index.end(), std::make_pair(htag[k], 0))->second;
  } else {
    for (int k = 0; k < nlocal; k++) rows[k] = k;    // several ranks: the host arrays are in device order
  }
  if (force->newton) host_borders();
  if (dynamic_cast<B200PairStyle *>(force->pair)->b200_peratom(ctx, eflag, vflag, rows.data(), nlocal)) fail();
  if (force->bond && atom->molecular != Atom::ATOMIC)
    if (dynamic_cast<B200BondStyle *>(force->bond)->b200_peratom_bond(ctx, eflag, vflag, rows.data(), nlocal)) fail();
}

/* ----------------------------------------------------------------------
   Atom::num_bond / bond_atom / bond_type and Atom::special / nspecial of the owned atoms,
   re-keyed by tag (CSR) for b200md_set_topology
------------------------------------------------------------------------- */

void VerletB200::push_topology()
{
  const int nlocal = atom->nlocal;
  const bigint natoms = atom->natoms;
  std::vector<int> boff(natoms + 2, 0), soff(natoms + 2, 0);
  for (int i = 0; i < nlocal; i++) {
    boff[atom->tag[i] + 1] = atom->num_bond[i];
    soff[atom->tag[i] + 1] = atom->nspecial[i][2];
  }
  for (bigint t = 1; t <= natoms + 1; t++) { boff[t] += boff[t - 1]; soff[t] += soff[t - 1]; }
  std::vector<int> bpart(boff[natoms + 1] + 1), btype(boff[natoms + 1] + 1), stag(soff[natoms + 1] + 1), slev(soff[natoms + 1] + 1);
  for (int i = 0; i < nlocal; i++) {
    const int t = atom->tag[i];
    for (int m = 0; m < atom->num_bond[i]; m++) {
      bpart[boff[t] + m] = (int) atom->bond_atom[i][m];
      btype[boff[t] + m] = atom->bond_type[i][m];
    }
    const int *ns = atom->nspecial[i];
    for (int m = 0; m < ns[2]; m++) {
      stag[soff[t] + m] = (int) atom->special[i][m];
      slev[soff[t] + m] = m < ns[0] ? 1 : (m < ns[1] ? 2 : 3);
    }
  }
  if (b200md_set_topology(ctx, natoms, boff.data(), bpart.data(), btype.data(), soff.data(), stag.data(), slev.data(),
                          force->special_lj, force->special_coul))
    fail();
  if (force->bond && dynamic_cast<B200BondStyle *>(force->bond)->b200_push_bond(ctx)) fail();
}

/* ----------------------------------------------------------------------
   host copies of the owned atoms, in the device's (bin-sorted) order
------------------------------------------------------------------------- */

void VerletB200::sync_host()
{
  if (!host_stale) return;
  int nlocal = 0, nghost = 0;
  if (b200md_get_nlocal(ctx, &nlocal, &nghost)) fail();
  if (comm->nprocs == 1 && (int) host_index.size() == nlocal) {
    // one rank: the host keeps its own atom order (so every other per-atom array -- mask,
    // molecule, bonds, special -- stays attached); device rows are scattered back by tag
    hx.resize(3 * (size_t) nlocal); hv.resize(3 * (size_t) nlocal); hf.resize(3 * (size_t) nlocal);
    htype.resize(nlocal); htag.resize(nlocal); himage.resize(nlocal);
    if (nlocal && b200md_get_atoms(ctx, hx.data(), hv.data(), hf.data(), htype.data(), htag.data(), himage.data())) fail();
    for (int k = 0; k < nlocal; k++) {
      auto it = std::lower_bound(host_index.begin(), host_index.end(), std::make_pair(htag[k], 0));
      const int i = it->second;
      for (int d = 0; d < 3; d++) {
        atom->x[i][d] = hx[3 * (size_t) k + d];
        atom->v[i][d] = hv[3 * (size_t) k + d];
        atom->f[i][d] = hf[3 * (size_t) k + d];
      }
      atom->image[i] = himage[k];
    }
    atom->nghost = 0;
    host_stale = false;
    return;
  }
  while (nlocal > atom->nmax) atom->avec->grow(0);
  atom->nlocal = nlocal;
  atom->nghost = 0;
  if (nlocal) {
    if (b200md_get_atoms(ctx, &atom->x[0][0], &atom->v[0][0], &atom->f[0][0], atom->type, atom->tag,
                         atom->image))
      fail();
  }
  int *mask = atom->mask;
  if (mask_by_tag.empty()) {
    const int m = (nlocal && comm->nprocs == 1) ? mask[0] : 1;    // uniform: keep; group all = bit 0
    for (int i = 0; i < nlocal; i++) mask[i] = m;
  } else {
    if (comm->nprocs > 1)
      error->all(FLERR, "Run style verlet/b200: non-uniform group masks need a single rank");
    for (int i = 0; i < nlocal; i++) {
      auto it = std::lower_bound(mask_by_tag.begin(), mask_by_tag.end(), std::make_pair((int) atom->tag[i], 0));
      mask[i] = it->second;
    }
  }
  if (atom->map_style != Atom::MAP_NONE) {
    atom->map_init();
    atom->map_set();
  }
  host_stale = false;
}

/* ----------------------------------------------------------------------
   run for N steps: device chunks that end on the host's next output step
------------------------------------------------------------------------- */

void VerletB200::run(int n)
{
  int done = 0;
  while (done < n) {
    if (timer->check_timeout(done)) {
      update->nsteps = done;
      break;
    }
    bigint togo = output->next - update->ntimestep;
    int chunk = n - done;
    if (togo > 0 && togo < (bigint) chunk) chunk = (int) togo;
    if (modify->n_end_of_step)    // next step an end_of_step fix is due (Modify::end_of_step: ntimestep % nevery == 0)
      for (int i = 0; i < modify->nfix; i++)
        if ((modify->fmask[i] & FixConst::END_OF_STEP) && modify->fix[i]->nevery > 0) {
          const bigint every = modify->fix[i]->nevery;
          const bigint due = (update->ntimestep / every + 1) * every - update->ntimestep;
          if (due < (bigint) chunk) chunk = (int) due;
        }

    for (const auto &fix : modify->get_fix_list())
      if (auto dev = dynamic_cast<B200FixStyle *>(fix))
        if (dev->b200_begin_chunk(ctx)) fail();
    timer->stamp();
    int nth = 0;
    if (b200md_run(ctx, chunk, update->dt, 0, nullptr, 0, &nth)) fail();
    timer->stamp(Timer::PAIR);    // device time of the whole chunk (pair + neigh + comm + nve)
    update->ntimestep += chunk;
    done += chunk;
    host_stale = true;

    long long nb = 0, nd = 0;
    b200md_neigh_stats(ctx, &nb, &nd, nullptr, nullptr);    // counters only: no list download
    neighbor->ncalls = (bigint) (nb - builds_at_setup);
    neighbor->ndanger = (bigint) (nd - danger_at_setup);

    bool eos = false;
    if (modify->n_end_of_step)
      for (int i = 0; i < modify->nfix; i++)
        if ((modify->fmask[i] & FixConst::END_OF_STEP) && modify->fix[i]->nevery > 0 &&
            update->ntimestep % modify->fix[i]->nevery == 0)
          eos = true;
    const bool out = update->ntimestep == output->next;
    if (out || eos) {
      // the last step of a chunk is evaluated with energy/virial tallies (Integrate::ev_set)
      ev_set(update->ntimestep);
      sync_host();
      publish_lists();
      publish_peratom();
      publish_tallies();
    }
    if (eos) {
      timer->stamp();
      modify->end_of_step();
      timer->stamp(Timer::MODIFY);
    }
    if (out) {
      timer->stamp();
      output->write(update->ntimestep);
      timer->stamp(Timer::OUTPUT);
    }
  }
}

/* ---------------------------------------------------------------------- */

void VerletB200::cleanup()
{
  sync_host();
  Verlet::cleanup();
}
