// ctx.h — internal device state of libb200md (not part of the ABI).
//
// Data layout in HBM (all FP64, int32 indices):
//   xq[nmax]      double4 {x,y,z,type-bits}: one 32-byte sector per atom, so a neighbor
//                 gather costs exactly one sector (the reference's AoS `double x[n][3]` +
//                 `int type[n]`, src/atom.h, straddles sectors; lib/gpu's numtyp4 has the same
//                 idea in FP32, lib/gpu/lal_lj.cu:16-25)
//   v[3][nmax], f[3][nmax], xhold[3][nmax]   SoA doubles (streamed only)
//   tag[nmax], image[nmax]                   int32
//   owned atoms 0..nlocal-1 are kept in neighbor-bin order (re-sorted at every rebuild),
//   ghosts nlocal..nall-1 in swap order like CommBrick::borders.
//   neighbor list: warp-tile transposed.  TPA lanes cooperate on one atom, A=32/TPA atoms
//   share a warp; tile t = atoms [t*A,(t+1)*A); slot k of atom i lives at
//       neigh[(size_t)t*pitch*A + (k/TPA)*32 + (i%A)*TPA + (k%TPA)]
//   so iteration `it` of a warp reads 128 contiguous bytes.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/b200md.h"

struct NcclState;  // comm.cu

struct EamTables {
  int ntypes = 0, nrho = 0, nr = 0, nfrho = 0, nrhor = 0, nz2r = 0;
  double rdr = 0, rdrho = 0, rhomax = 0, cutforcesq = 0;
  // repacked tables, rows 0..n (row m as in the reference, 1-based)
  double *frho = nullptr;   // [nfrho][nrho+1][8]  c0..c6,pad (embedding: once per atom)
  double *rhoA = nullptr;   // [nrhor][nr+1][4]    density value cubic c3..c6
  double *z2rV = nullptr;   // [nz2r][nr+1][4]     pair-term value cubic c3..c6
  double *tabC = nullptr;   // [nr+1][8]           one-element runs: {rhoA row, z2rV row}
  double4 *xfp = nullptr;   // [nmax]              one-element runs: {x,y,z,fp} gather record
  int xfp_cap = 0;
  unsigned *pmask = nullptr; // [tiles][4][32] half list: bit `it` of a lane's 128 = its it-th list entry is inside the force cutoff this step
  size_t pmask_cap = 0;     //              (written by the density pass, walked by the force pass)
  bool one = false;
  double scale11 = 1.0;
  int *type2frho = nullptr; // [ntypes+1]
  int *type2rhor = nullptr; // [(ntypes+1)^2]
  int *type2z2r = nullptr;  // [(ntypes+1)^2]
  double *scale = nullptr;  // [(ntypes+1)^2]
};

// kernel-argument view of EamTables (pair_eam.cu, peratom.cu)
struct EamParams {
  int ntypes, nrho, nr;
  double rdr, rdrho, rhomax, cutforcesq, scale11;
  const double *frho, *rhoA, *z2rV, *tabC, *scale;
  const int *type2frho, *type2rhor, *type2z2r;
};

// molecular topology keyed by atom tag (bond.cu); CSR offsets are [natoms+2], row t = tag t
struct Topology {
  bool molecular = false;
  long long natoms = 0;
  int *bond_off = nullptr, *bond_partner = nullptr, *bond_type = nullptr;  // bonds stored on tag t (newton_bond on)
  int *spec_off = nullptr, *spec_tag = nullptr, *spec_which = nullptr;     // special neighbors of tag t: tag, 1/2/3
  int special_flag[4] = {1, 1, 1, 1};
  double special_lj[4] = {1, 1, 1, 1};
  int *map = nullptr;            // [natoms+1] tag -> local index (owned copy preferred), -1 if absent
  int nbondtypes = 0;
  double *fene = nullptr;        // [nbondtypes+1][4] k, r0, epsilon, sigma
  double eng_bond = 0, virial_bond[6] = {0, 0, 0, 0, 0, 0};
};

// fix langevin + the device RanMars stream (langevin.cu)
struct Langevin {
  bool on = false;
  double t_start = 0, t_stop = 0, t_period = 0, dt = 0;
  std::vector<double> ratio;       // [ntypes+1]
  bool window = false;             // begin/end/now set by the caller (b200md_fix_langevin_window)
  long long begin = 0, end = 0, now = 0;
  uint32_t *state[2] = {nullptr, nullptr};  // device [97]: generator state the chunk layout starts from
  uint32_t h_state[B200MD_RANMARS_LAG];     // host copy of a known generator state ...
  long long h_position = 0;                 // ... and the number of draws made before it
  long long position = 0;          // draws made so far
  uint32_t *jumpT = nullptr;       // device [24][97][97] transposed jump tables (chunk layout)
  uint32_t *stepT = nullptr;       // device [97][97] transposed jump by one evaluation's draws
  uint32_t *chunk_state = nullptr; // device [chunk_cap][97] resident generator state of every chunk
  int chunk_cap = 0, stride = 0;   // stride: draws per evaluation the layout was made for
  long long nevals = 0;
  uint32_t *rnd = nullptr;         // device [rnd_cap] this evaluation's draws, 24-bit integers
  int rnd_cap = 0;
  int *rank_by_tag = nullptr;      // device [maxtag+1] host-order index of a tag
  int nrank = 0;
  double *gfac = nullptr;          // device [ntypes+1][2] gfactor1, gfactor2
  int gfac_n = 0;
  // Atom::sort schedule of the host the stream order mirrors (atom.cpp:102-103, 2252; verlet.cpp:285)
  long long sortfreq = 1000, nextsort = 0;
  bool resort_pending = false;     // set by b200md_run for the rebuild that has to re-rank first
  unsigned long long *sort_keys = nullptr;   // langevin_resort scratch: [2][sort_cap] keys, values, radix-sort workspace
  int *sort_vals = nullptr;
  void *sort_tmp = nullptr;
  int sort_cap = 0;
  size_t sort_tmp_bytes = 0;
};

struct Swap {
  int dim, ineed, sendflag, sendproc, recvproc;
  int pbc[3];
  double slablo, slabhi;
  int sendnum = 0, recvnum = 0, firstrecv = 0;
  int *sendlist = nullptr;  // device
  int sendcap = 0;
};

// "direct" halo (comm.cu, B200MD_HALO=direct): after the staged borders have placed the ghosts, every
// ghost knows its OWNER (rank, index on that rank, accumulated periodic shift), so the per-step
// traffic no longer has to be forwarded dimension by dimension: one pack, ONE NCCL group with a
// message per peer rank, one unpack — instead of three dependent pack/group stages.
struct DirectHalo {
  bool on = false, ready = false;
  int nreq = 0, nreq_remote = 0, nserve = 0;
  std::vector<int> req_cnt, req_off, serve_cnt, serve_off;  // per rank; my requests to / the requests I serve for rank p
  int *req_idx = nullptr;     // device [nreq]   index on the owner, grouped by owner rank (self last)
  int *ghost_slot = nullptr;  // device [nreq]   my ghost that request fills
  int *serve_idx = nullptr;   // device [nserve] my owned atoms other ranks asked for, grouped by asking rank
  int req_cap = 0, serve_cap = 0;
  int *d_cnt = nullptr;       // device [nranks] + [nranks] cursor + [nranks*nranks] gathered matrix
  int *h_cnt = nullptr;       // pinned, same layout
};

// "direct" halo on several ranks (comm.cu, option halo=direct, the default): every owned atom is classified ONCE per
// rebuild by the up to 26 directions (dx,dy,dz) it has to appear in as a ghost, and by the one direction it migrates
// in; a direction maps to one peer rank and one periodic shift.  Owner and holder then talk directly: one message per
// peer rank per exchange, no forwarding stages, no owner bookkeeping, no all-gather.  Rebuild-time messages have a
// capacity both sides derive from the previous rebuild's counts and carry their true counts in a header, so a
// rebuild costs one NCCL group + one host sync for the migration and one + one for the ghosts.
#define H26_NDIR 27
#define H26_HDR 32   // ints per message header
struct Halo26 {
  bool on = false, ready = false;
  int peer[H26_NDIR];           // rank a copy sent in direction d lands on; -1 = direction unused
  int shift[H26_NDIR][3];       // periodic shift multiples applied by the sender
  int npeers = 0;               // distinct REMOTE peers
  std::vector<int> peers;       // ascending rank
  std::vector<std::vector<int>> dirs_to, dirs_from;   // per remote peer: directions I send in / receive in, ascending
  std::vector<int> dirs_self;   // directions that land on this rank (a dimension with one rank)
  int cap_b[H26_NDIR], cap_m[H26_NDIR];     // capacities (atoms) of my outgoing border / migrant segments
  int capin_b[H26_NDIR], capin_m[H26_NDIR]; // ... of the incoming ones (= the sender's cap for that direction)
  int prev_b[H26_NDIR], prevb_in[H26_NDIR]; // counts of the previous rebuild (what the next capacities derive from)
  bool have_prev = false;
  int *d_cnt = nullptr;         // device [64]: border counters [0,27), migrant counters [32,59)
  int *h_cnt = nullptr;         // pinned [64]
  int *d_hdr_send = nullptr, *d_hdr_recv = nullptr;   // device [npeers][H26_HDR]
  int *h_hdr = nullptr;         // pinned [npeers][H26_HDR]
  int *list = nullptr;          // device: border candidates of direction d at [lo_b[d], lo_b[d] + cap_b[d])
  int list_cap = 0;
  int lo_b[H26_NDIR + 1];
  int *dead = nullptr;          // device [nmax]: 1 = atom left this rank (dropped by the next sort)
  int dead_cap = 0;
  int *send_idx = nullptr;      // device [nsend]: owned atom of every outgoing ghost copy, remote peers first (by peer,
  int *send_dir = nullptr;      //   direction), then the copies that stay on this rank; its direction
  int send_cap = 0;
  int nsend = 0, nsend_remote = 0, nghost_remote = 0;
  std::vector<int> send_off, send_cnt, recv_off, recv_cnt;   // per remote peer, in atoms (per-step message sizes)
  char *sbuf = nullptr, *rbuf = nullptr;   // rebuild-time message regions
  size_t sbuf_cap = 0, rbuf_cap = 0;
  int ndead = 0;                // leavers of the last exchange, not yet dropped
  long long nredo = 0;          // rebuilds that had to repeat an exchange because a capacity was exceeded
};

// option graph (api.cu): the non-rebuild steps between two neighbor-list builds replayed as ONE CUDA graph launch — a
// WHILE conditional node whose body is a whole timestep and whose condition is Neighbor::decide evaluated on the device
struct StepGraph {
  int mode = -1;                 // -1 auto (one rank, small systems), 0 off, 1 on
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  long long built_for = -1;      // c->nbuilds the body was captured for (its kernel arguments change with every list)
  double dtv = 0, dtf = 0;
  int launches_per_body = 0, pair_launches_per_body = 0;
  int *d_loop = nullptr, *h_loop = nullptr;   // [0] bodies left, [1] ago, [2] bodies done, [3] rebuild decided for the next step
  long long nlaunch = 0, nsteps = 0;
};

struct b200md_ctx {
  int device = 0, rank = 0, nranks = 1;
  cudaStream_t stream = nullptr;
  // units
  double boltz = 1, mvv2e = 1, ftm2v = 1, nktv2p = 1;
  // box
  double boxlo[3], boxhi[3], prd[3], sublo[3], subhi[3];
  int periodic[3] = {1, 1, 1}, procgrid[3] = {1, 1, 1}, myloc[3] = {0, 0, 0};
  int procneigh[3][2];
  bool box_set = false;
  // atoms
  int nlocal = 0, nghost = 0, nmax = 0, ntypes = 0;
  double natoms_global = 0;
  double4 *xq = nullptr, *xq_tmp = nullptr;
  double *v = nullptr, *v_tmp = nullptr;  // 3*nmax each, SoA with stride nmax
  double *f = nullptr;
  double *xhold = nullptr;
  int *tag = nullptr, *tag_tmp = nullptr, *image = nullptr, *image_tmp = nullptr;
  double *mass = nullptr;  // device [ntypes+1]
  std::vector<double> h_mass;
  // neighbor config
  int list_kind = B200MD_LIST_FULL;
  double *cutneighsq = nullptr;  // device (ntypes+1)^2
  std::vector<double> h_cutneighsq;
  double cutneighmax = 0, skin = 0, triggersq = 0, cutghost = 0;
  int every = 1, delay = 0, check = 1;
  int ago = 0;
  long long nbuilds = 0, ndanger = 0;
  bool neigh_configured = false;
  b200md_bin_geometry bin;
  int mbins = 0;
  // binning scratch
  int *atom_bin = nullptr;    // [nmax]
  int *cell_count = nullptr;  // [mbins+1]
  int *cell_start = nullptr;  // [mbins+1]
  int *cell_fill = nullptr;   // [mbins]
  int *cell_atoms = nullptr;  // [nmax]  atom indices in cell order
  double4 *cell_xq = nullptr; // [nmax]  positions in cell order
  int *perm = nullptr;        // [nmax]
  int cell_cap = 0;
  int *scan_block = nullptr;  // scan partials
  int scan_cap = 0;
  // stencil rows: (dy,dz) rows with x half-extent
  int nrows = 0;
  int *row_off = nullptr;  // device [nrows] flattened (dz*mbiny+dy)*mbinx
  int *row_sx = nullptr;   // device [nrows]
  int *row_flag = nullptr; // device [nrows] half-stencil role (0 skip,1 full row,2 own row)
  // neighbor list
  int tpa = 2;        // lanes per atom in the force kernels (1..32); measured best on B200
  int tpa_build = 1;  // lanes per atom in the list build (1..32); measured best on B200 for lj, eam and chain
  bool tpa_build_user = false;
  bool tpa_user = false;  // set by b200md_set_lanes_per_atom; otherwise the pair style picks (lj 2, eam 1)
  int use_cluster = -1;     // option cluster: on (1) / off (0) / auto (-1: lj/cut yes; eam no, its table look-ups are
                            // per pair either way and the masked-out slots of a cluster would fetch rows for nothing)
  bool cluster = false;     // layout of the current list: units are clusters of 4 atoms, entries carry a 4-bit pair mask
  int pitch = 0;  // max neighbors per list unit (atom or cluster; multiple of 32)
  int *neigh = nullptr;
  size_t neigh_cap = 0;
  int *numneigh = nullptr;  // [nmax]
  // tiles (warps) split by 'has a ghost neighbor' for compute/halo overlap
  int *tile_ghost = nullptr, *tile_pos = nullptr, *tiles_int = nullptr, *tiles_bnd = nullptr;
  int tile_cap = 0, n_tiles_int = 0, n_tiles_bnd = 0;
  cudaStream_t comm_stream = nullptr, bnd_stream = nullptr;
  cudaEvent_t ev_x = nullptr, ev_halo = nullptr, ev_fp = nullptr, ev_fphalo = nullptr;
  cudaEvent_t ev_bnd = nullptr, ev_rev = nullptr, ev_rb = nullptr, ev_rr = nullptr, ev_clr = nullptr;  // half-list overlap
  double split = 0.5;  // share of the interior tiles evaluated before the boundary tiles (half-list overlap)
  bool overlap = true;
  bool fuse_nve = true;   // option fuse_nve: final_integrate(n) + initial_integrate(n+1) in one pass (nve.cu), bit-identical
  bool cell_by_x = true;   // option cell_order: atoms of a bin ordered by x (then index / tag) instead of index / tag
  double bin_scale = 1.0;  // neighbor bins of bin_scale x cutneighmax/2 in use (neigh_choose_layout)
  double bin_scale_user = 0; // option bin_scale: > 0 fixes it, 0 = auto
  bool eam_mask = true;   // option eam_mask: the force pass walks only the pairs the density pass found inside the cutoff
  int eam_smem = -1;      // option eam_smem: one-element EAM tables in shared memory (swizzled rows): -1 auto = the force
                          // pass only (measured: force 815 -> 756 us, density pass unchanged), 1 every pass, 0 never
  bool phase_timing = false;  // option phase_timing: per-phase CUDA events in b200md_run
  bool trace = false;         // option trace: timestamps of both streams' milestones of every step, printed to stderr after b200md_run
  std::vector<std::pair<std::string, cudaEvent_t>> trace_ev;
  // comm
  b200md_swap_schedule sched;
  std::vector<Swap> swaps;
  NcclState *nccl = nullptr;
  int *ghost_src = nullptr;       // single-rank fast path: flattened owner index per ghost
  int *ghost_pbc = nullptr;       // packed periodic-shift multiples (px+8)|(py+8)<<4|(pz+8)<<8
  int *ghost_rank = nullptr;      // owner rank per ghost (direct halo)
  DirectHalo direct;
  Halo26 h26;
  StepGraph sg;
  int halo_mode = 2;              // option halo: 0 staged, 1 owner (staged borders that track owners), 2 direct (halo26.cu)
  bool all_self = true;           // every swap partner is this rank -> flattened forward path
  bool self26 = true;             // option self_border: one rank, one ghost layer: all borders in one pass (comm.cu)
  int *s26_cnt = nullptr, *s26_hcnt = nullptr;   // its per-direction totals (device / pinned host)
  double *sendbuf = nullptr, *recvbuf = nullptr;
  size_t sendcap = 0, recvcap = 0;
  // pair
  int pair_style = B200MD_PAIR_NONE;
  double *ljtab = nullptr;  // device (ntypes+1)^2 * 8: cutsq,lj1,lj2,lj3,lj4,offset,pad,pad
  std::vector<double> h_ljtab;
  double cutforcemax = 0;
  EamTables eam;
  Topology topo;
  Langevin lang;
  bool host_order = false;   // owned atoms are still in the order of the last b200md_set_atoms
  double *rho = nullptr, *fp = nullptr;  // [nmax]
  double *peratom = nullptr;             // [7][peratom_cap] eatom, vatom xx yy zz xy xz yz (peratom.cu; output steps only)
  int peratom_cap = 0;
  // reductions
  double *partials = nullptr;  // device [npartial_blocks*8]
  int partial_cap = 0;
  double *d_result = nullptr;  // device [16]
  double *h_result = nullptr;  // pinned [16]
  int *d_flags = nullptr;      // device [8]: 0 rebuild flag, 1 overflow, 2 nonfinite, 3 counter
  int *h_flags = nullptr;      // pinned [8]
  cudaStream_t copy_stream = nullptr;   // b200md_get_atoms_async: D2H copies next to the next step's upload and kernels
  cudaEvent_t ev_packed = nullptr, ev_dl_x = nullptr, ev_dl_all = nullptr;
  char *dl_stage = nullptr;
  size_t dl_cap = 0;
  bool dl_pending = false;
  std::vector<cudaEvent_t> ev_dl_part;  // x download in pieces (b200md_wait_download_part): one event per piece
  int dl_parts = 0, dl_n = 0;           // pieces and atoms of the download in flight
  char *stage = nullptr;  // device staging for host<->device atom transfers
  size_t stage_cap = 0;
  // last energies
  double eng_vdwl = 0, virial[6] = {0, 0, 0, 0, 0, 0};
  bool ev_valid = false;
  long long step = 0;
  // timing
  b200md_timing timing;
  long long launches = 0, pair_launches = 0;
  std::vector<cudaEvent_t> ev_pool;
};

// ---- error helpers (api.cu) ----
void b200md_set_error(const std::string &msg);
#define CUDA_TRY(call)                                                                   \
  do {                                                                                   \
    cudaError_t _e = (call);                                                             \
    if (_e != cudaSuccess) {                                                             \
      b200md_set_error(std::string("b200md: CUDA error: ") + cudaGetErrorString(_e) +    \
                       " at " + __FILE__ + ":" + std::to_string(__LINE__));              \
      return B200MD_ERR;                                                                 \
    }                                                                                    \
  } while (0)
#define TRY(call)                      \
  do {                                 \
    int _r = (call);                   \
    if (_r != B200MD_OK) return _r;    \
  } while (0)

// list units: what a group of `tpa` lanes works on — an atom, or a cluster of 4 atoms (neigh.cu)
#ifndef CL
#define CL 4   // atoms per i-cluster of the cluster-pair list (util.cuh)
#endif
static inline int list_units(const b200md_ctx *c) { return c->cluster ? (c->nlocal + CL - 1) / CL : c->nlocal; }
static inline int list_tiles(const b200md_ctx *c) { const int A = 32 / c->tpa; return (list_units(c) + A - 1) / A; }

// ---- kernels' host launchers ----
// neigh.cu
int neigh_grow_atoms(b200md_ctx *c, int need);   // grow per-atom arrays to >= need
int neigh_setup_bins(b200md_ctx *c);             // geometry + stencil rows
int neigh_choose_layout(b200md_ctx *c);          // bin size for the configured pair style / topology (every rebuild)
int neigh_pbc(b200md_ctx *c);                    // Domain::pbc on owned atoms
int neigh_sort_owned(b200md_ctx *c);             // bin-order sort of owned atoms
int neigh_bin_all(b200md_ctx *c);                // cell lists over owned+ghost
int neigh_build_list(b200md_ctx *c);             // NPair build
int neigh_save_xhold(b200md_ctx *c);
int dev_exclusive_scan(b200md_ctx *c, const int *in, int *out, int n);
// comm.cu
int comm_setup(b200md_ctx *c);
int comm_exchange(b200md_ctx *c);
int comm_borders(b200md_ctx *c);
int comm_forward(b200md_ctx *c);
int comm_forward_scalar(b200md_ctx *c, double *arr);  // EAM fp
int comm_reverse_force(b200md_ctx *c);
int comm_reverse_scalar(b200md_ctx *c, double *arr);  // EAM rho (half lists)
int comm_reverse_rho(b200md_ctx *c);
int comm_reverse_column(b200md_ctx *c, double *arr);  // like _scalar, also on full lists (per-atom bond tallies)
int comm_allreduce_sum(b200md_ctx *c, double *hostvals, int n);
int comm_allreduce_max_int(b200md_ctx *c, int *hostval);
void comm_free(b200md_ctx *c);
// pair_lj.cu / pair_eam.cu
int pair_lj_compute(b200md_ctx *c, int eflag, int vflag);
int pair_lj_launch(b200md_ctx *c, bool ev, const int *tiles, int ntl);
int pair_eam_launch(b200md_ctx *c, bool ev, int pass, const int *tiles, int ntl);
int pair_eam_fp_comm(b200md_ctx *c);
int pair_eam_embed(b200md_ctx *c, bool ev, double *e_embed);  // half-list path: F(rho), F'(rho) of owned atoms
int pair_eam_compute(b200md_ctx *c, int eflag, int vflag);
EamParams pair_eam_params(const b200md_ctx *c);
// peratom.cu
int peratom_compute(b200md_ctx *c, int eflag_atom, int vflag_atom, int which);
void peratom_free(b200md_ctx *c);
// bond.cu
void topo_free(b200md_ctx *c);
int topo_set(b200md_ctx *c, long long natoms, const int *bond_off, const int *bond_partner, const int *bond_type,
             const int *spec_off, const int *spec_tag, const int *spec_which, const double special_lj[4],
             const double special_coul[4]);
int bond_fene_set(b200md_ctx *c, int nbondtypes, const double *k, const double *r0, const double *eps, const double *sigma);
int topo_build_map(b200md_ctx *c);
int bond_compute(b200md_ctx *c, int eflag, int vflag);
// langevin.cu
int langevin_config(b200md_ctx *c, double t_start, double t_stop, double t_period, int seed, const double *ratio, double dt);
int langevin_set_dt(b200md_ctx *c, double dt);
int langevin_rank_from_host_order(b200md_ctx *c);
int langevin_post_force(b200md_ctx *c, bool with_final, double dtf, bool with_next = false, double dtv = 0.0,
                        bool check_dist = false);
int langevin_resort(b200md_ctx *c);
int langevin_test_stream(b200md_ctx *c, int seed, int n, int fills, double *out);
void langevin_free(b200md_ctx *c);
// nve.cu
int nve_initial(b200md_ctx *c, double dtv, double dtf, bool check_dist);
int nve_final(b200md_ctx *c, double dtf);
int nve_final_initial(b200md_ctx *c, double dtv, double dtf, bool check_dist);
int thermo_ke(b200md_ctx *c, double *sum_mv2);
int reduce_partials(b200md_ctx *c, int nblocks, int nvals, double *host_out);
int ensure_partials(b200md_ctx *c, int nblocks);
