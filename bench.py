#!/usr/bin/env python3
"""bench.py — atom-steps/s of the short-range MD hot path (neighbor build -> lj/cut | eam -> fix nve)
on N B200s of one node, measured on the reference's own bench inputs.

  python bench.py --gpus N --steps K --warmup W [--impl reference]

A "step" is one MD timestep of the workload (Verlet::run body, /root/reference/src/verlet.cpp:246-359)
including the neighbor rebuilds that fall inside the K steps.  One JSON line on rank 0:

  value / e2e / roofline   bench/in.lj scaled to 2 048 000 atoms per GPU (-var x/y/z = 4,4,4 / 8,4,4 / 8,8,4 /
                           8,8,8: config C4, weak scaling), one process per GPU, spatial decomposition
  parity_check             BEFORE anything is timed: bench/in.lj and bench/in.eam at 32 000 atoms, 100 steps on
                           the same N ranks with the halo mode and list kind of the timed regions; thermo at steps
                           0/50/100 must equal the reference's own logs (bench/log.6Oct16.{lj,eam}.fixed.icc.1) in
                           all 8 printed digits.  A mismatch prints the offending line and exits non-zero.
  secondary.eam            bench/in.eam (Cu_u3) at 2 048 000 atoms per GPU (weak), with its own roofline
  secondary.eam_alloy_strong   config C5: in.eam with eam/alloy Cu_mishin1, 4 000 000 atoms in TOTAL on the N ranks
                           (strong scaling, atom migration at every rebuild); its thermo at steps 0/50/100 is checked
                           against tests/golden/c5_thermo.json (the unmodified reference, 4 M atoms)
  secondary.lj_32k / eam_32k   the 32 000-atom configs C1 / C2 as shipped (latency regime)
  secondary.chain          config C3: bench/in.chain.scaled at 2 048 000 beads per GPU (parity_check.chain: 32 000 beads against
                           tests/golden/chain.npz from the unmodified reference)
  cpu_baseline (N=1)       the reference's own CPU implementation on this box's host cores, bounded sample
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SCALE = {1: (4, 4, 4), 2: (8, 4, 4), 4: (8, 8, 4), 8: (8, 8, 8)}
# algorithmic bytes: force pass(es) per owned atom per step, and whole step per atom-step.
# N = full-list length, N_h = N/2.  full list: LJ 28N+52, EAM 64N+104; half list: LJ 52 N_h + 52 (4 index + 24 x_j
# gathered + 24 f_j RED payload per listed pair), EAM (36 + 60 r) N_h + 136 with r = in-cutoff share (below).
N_FULL = {"lj": 75.2, "eam": 75.5, "c5": 88.8}
BYTES_KERNEL = {k: {"full": (28 if k == "lj" else 64) * n + (52 if k == "lj" else 104),
                    "half": (52 if k == "lj" else 96) * n / 2 + (52 if k == "lj" else 104)} for k, n in N_FULL.items()}
# eam half list with in-cutoff masks: the density pass moves 4 + 24 + 8 (RED) per LISTED pair, the force pass
# 4 + 24 + 8 + 24 (RED) per pair INSIDE the force cutoff only — share (cutforce / cutneigh)^3 of a list built with skin 1.0 —
# plus 32 bytes of masks per atom written and read
IN_CUTOFF = {"eam": (4.95 / 5.95) ** 3, "c5": (5.50679 / 6.50679) ** 3}
for _k, _r in IN_CUTOFF.items():
    BYTES_KERNEL[_k]["half"] = (36 + 60 * _r) * N_FULL[_k] / 2 + 104 + 32
BYTES_STEP = {"lj": 2480.0, "eam": 5510.0, "c5": 6420.0}
KERNELS = {("lj", "half"): "k_lj<half>", ("lj", "full"): "k_lj<full>",
           ("eam", "half"): "k_eam_rho_half<mask>+k_eam_embed+k_eam_force<half,mask>", ("eam", "full"): "k_eam_rho_fp+k_eam_force",
           ("c5", "half"): "k_eam_rho_half<mask>+k_eam_embed+k_eam_force<half,mask>", ("c5", "full"): "k_eam_rho_fp+k_eam_force"}
# thermo printed by the reference's own logs (bench/log.6Oct16.{lj,eam}.fixed.icc.1): T, E_pair, TotEng, Press
LOG = {"lj": {0: (1.44, -6.7733681, -4.6134356, -5.0197073), 100: (0.7574531, -5.7585055, -4.6223613, 0.20726105)},
       "eam": {0: (1600.0, -113280.0, -106662.09, 18703.573), 50: (781.69049, -109873.35, -106640.13, 52273.088),
               100: (801.832, -109957.3, -106640.77, 51322.821)}}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (via NVML)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if not self.nv:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.01)

    def result(self):
        self.stop_flag = True
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}
        return {"sm_mhz": float(np.median(self.samples)), "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------------------------------
# workloads
# ------------------------------------------------------------------------------------------------------------
def workload_xyz(wl, ngpus):
    if wl in ("lj", "eam"):
        return SCALE.get(ngpus, (4 * ngpus, 4, 4))
    if wl == "c5":
        return (5, 5, 5)
    return (1, 1, 1)       # lj_32k, eam_32k


def workload_desc(wl, ngpus):
    xyz = workload_xyz(wl, ngpus)
    natoms = 4 * 8000 * xyz[0] * xyz[1] * xyz[2]
    kind = {"lj": "lj", "lj_32k": "lj", "eam": "eam", "eam_32k": "eam", "c5": "c5"}[wl]
    text = {"lj": "bench/in.lj -var x %d -var y %d -var z %d (%d atoms, fcc lattice, lj/cut 2.5, neigh every 20, NVE)",
            "eam": "bench/in.eam -var x %d -var y %d -var z %d (%d atoms, fcc lattice, eam Cu_u3, neigh check yes, NVE)",
            "c5": "bench/in.eam with pair_style eam/alloy Cu_mishin1 -var x %d -var y %d -var z %d (%d atoms in total, "
                  "strong scaling, neigh check yes, NVE)"}[kind] % (xyz[0], xyz[1], xyz[2], natoms)
    return kind, xyz, natoms, text


def ref_script(kind, xyz, warm, steps, thermo=0):
    from oracle import oracle as O
    if kind == "lj":
        body = O.IN_LJ.format(nx=20 * xyz[0], ny=20 * xyz[1], nz=20 * xyz[2])
    elif kind == "eam":
        body = O.IN_EAM.format(nx=20 * xyz[0], ny=20 * xyz[1], nz=20 * xyz[2], style="eam",
                               coeff="1 1 " + os.path.join(O.POT_DIR, "Cu_u3.eam"))
    else:
        body = O.IN_EAM.format(nx=20 * xyz[0], ny=20 * xyz[1], nz=20 * xyz[2], style="eam/alloy",
                               coeff="* * " + os.path.join(O.POT_DIR, "Cu_mishin1.eam.alloy") + " Cu")
    body = body.replace("atom_modify map array\n", "")
    if thermo:
        body += "thermo %d\n" % thermo
    return body + ("run %d\n" % warm if warm > 0 else "") + "run %d\n" % steps


def run_reference(kind, xyz, warm, steps, cores):
    """the reference's own CPU implementation (oracle/_ref/lmp, OPENMP package on all host cores)"""
    import re
    from oracle import oracle as O
    _, atoms, nst, out = O.ref_exe_run(ref_script(kind, xyz, warm, steps), threads=cores, omp=True, timeout=3000)
    loops = re.findall(r"Loop time of ([0-9.eE+-]+) on \d+ procs for (\d+) steps with (\d+) atoms", out)
    lt, nst, atoms = float(loops[-1][0]), int(loops[-1][1]), int(loops[-1][2])
    return atoms * nst / lt, lt, atoms


T_START = time.time()


def progress(msg):
    """phase markers on stderr (rank 0), so a run that stalls says where"""
    if int(os.environ.get("RANK", "0")) == 0:
        print("[bench %6.1fs] %s" % (time.time() - T_START, msg), file=sys.stderr, flush=True)


def fmt8(v):
    return float("%.8g" % v)


def thermo_mismatch(recs, want, per):
    """recs: {step: thermo dict}; want: {step: (T, E_pair, TotEng, Press)} printed with 8 digits"""
    for step, w in want.items():
        t = recs.get(step)
        if t is None:
            return "no thermo record at step %d" % step
        got = (t["temp"], t["pe"] / per, (t["pe"] + t["ke"]) / per, t["press"])
        for name, a, b in zip(("T", "E_pair", "TotEng", "Press"), got, w):
            if abs(fmt8(a) - b) > 2e-8 * abs(b):
                return "step %d %s: got %.10g, the reference prints %.8g" % (step, name, a, b)
    return None


class Job:
    """one process per GPU; NCCL id shipped over torch.distributed"""

    def __init__(self):
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.dist = None
        self.affinity = None
        try:      # pinned host buffers are first touched by this process: keep it on the CPUs next to its GPU
            import pynvml
            pynvml.nvmlInit()
            h = pynvml.nvmlDeviceGetHandleByIndex(self.local_rank)
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
            cpus = [64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1]
            if cpus:
                os.sched_setaffinity(0, cpus)
                self.affinity = "%d cpus near gpu %d" % (len(cpus), self.local_rank)
        except Exception as e:
            self.affinity = "not set (%s)" % str(e)[:60]
        import b200pkg
        self.P = b200pkg.load()
        self.B = self.P.b200md
        if self.world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(self.local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
            self.dist = dist

    def open(self, options=()):
        B = self.B
        ctx = B.Context(self.local_rank, self.rank, self.world)
        if self.world > 1:
            import torch
            idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if self.rank == 0:
                idt = torch.frombuffer(bytearray(B.Context.unique_id()), dtype=torch.uint8).cuda()
            self.dist.broadcast(idt, 0)
            ctx.comm_init(bytes(idt.cpu().numpy().tobytes()))
        for k, v in options:
            ctx.set_option(k, v)
        return ctx

    def barrier(self):
        if self.dist is not None:
            import torch
            self.dist.barrier()
            torch.cuda.synchronize()

    def maxr(self, v):
        if self.dist is None:
            return v
        import torch
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def configure(self, ctx, kind, xyz, lk):
        B, P = self.B, self.P
        list_kind = B.LIST_HALF_NEWTON if lk == "half" else B.LIST_FULL
        if kind == "lj":
            return P.inputs.in_lj(ctx, *xyz, list_kind=list_kind)
        if kind == "eam":
            return P.inputs.in_eam(ctx, *xyz, list_kind=list_kind)
        return P.inputs.in_eam(ctx, *xyz, style="eam/alloy", list_kind=list_kind)


def measure(job, ctx, cfg, natoms, steps, warmup, sampler=None):
    """W warm-up steps, then K steps timed on the device (CUDA events on the context's stream around the K steps,
    max over ranks) between barriers; then 20 more steps with per-phase events for the kernel figures"""
    ctx.run(max(warmup, 3), cfg["dt"], 0)
    if sampler:
        sampler.start()
    job.barrier()
    t0 = time.perf_counter()
    ctx.run(steps, cfg["dt"], 0)
    job.barrier()
    wall_ms = job.maxr((time.perf_counter() - t0) * 1e3)
    tm = ctx.timing()
    clocks = sampler.result() if sampler else None
    dev_ms = job.maxr(tm["total_ms"])
    ctx.set_option("phase_timing", "on")
    ctx.run(20, cfg["dt"], 0)
    ctx.set_option("phase_timing", "off")
    tp = ctx.timing()
    phases = {k: job.maxr(tp[k + "_ms"]) / 20 for k in ("pair", "neigh", "comm", "modify")}
    return dict(value=natoms * steps / (dev_ms / 1e3), ms_per_step=dev_ms / steps, device_ms=dev_ms, wall_ms=wall_ms,
                launches=int(tm["launches"]), rebuilds=int(tm["nbuilds"]), phases=phases, clocks=clocks)


def roofline_of(kind, lk, m, nloc, ngpus, layout=None):
    hbm, which = peaks()
    bk = BYTES_KERNEL[kind][lk]
    kname = KERNELS[(kind, lk)]
    if layout and layout["cluster"]:
        # cluster-pair list: what the kernel's algorithm moves per owned atom and launch is E stored entries
        # x (4 index + 32 x_j record [+ 24 f_j RED payload on the half list]) + 32 x_i + 24 f_i, E measured from the list
        e_per_atom = layout["entries"] / max(nloc, 1)
        bk = e_per_atom * (36 + (24 if lk == "half" else 0)) + 56
        kname = "k_lj_cl<%s> (cluster-pair list, %.1f entries/atom, %d lanes/cluster)" % (lk, e_per_atom, layout["lanes"])
    pair_ms = m["phases"]["pair"]
    achieved = bk * nloc / (pair_ms / 1e3) / 1e9
    r = {"bound": "hbm", "kernel": kname, "list": lk, "achieved": achieved, "peak": hbm, "peak_source": which,
         "unit": "GB/s", "frac": achieved / hbm, "traffic": None, "kernel_ms": pair_ms, "bytes_per_atom_launch": bk,
         "step_achieved": BYTES_STEP[kind] * (m["value"] / ngpus) / 1e9,
         "step_frac": BYTES_STEP[kind] * (m["value"] / ngpus) / 1e9 / hbm,
         "phases_ms_per_step": m["phases"],
         "note": "algorithmic gather bytes (every listed neighbor record counted as an HBM read) over the measured copy "
                 "bandwidth; atoms are bin-sorted, so most gathers hit L1/L2 and the unit that binds is the L1 LSU data "
                 "pipe, see binding_unit / DESIGN.md §3"}
    # DRAM traffic and the binding unit of THIS kernel variant from its `ncu --set full` capture (profiles/), if one is
    # committed for exactly this kernel name
    tr = os.path.join(ROOT, "profiles", "traffic_%s_%s.json" % (kind, lk))
    if os.path.exists(tr):
        t = json.load(open(tr))
        if t.get("kernel_variant") == kernel_variant():
            r["traffic"] = t.get("dram_bytes_per_launch")
            r["binding_unit"] = t.get("binding_unit")
    return r


def kernel_variant():
    p = os.path.join(ROOT, "profiles", "KERNEL_VARIANT")
    return open(p).read().strip() if os.path.exists(p) else None


def main():
    wd = int(os.environ.get("B200MD_BENCH_WATCHDOG", "1500"))
    if wd > 0:      # a rank that fails leaves its peers waiting in NCCL: dump the Python stacks and exit instead of hanging
        import faulthandler
        faulthandler.dump_traceback_later(wd, exit=True)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="lj", choices=["lj", "eam", "c5"],
                    help="the workload reported as `value` (default lj = config C4; the others are always under `secondary`)")
    ap.add_argument("--lanes", type=int, default=0)
    ap.add_argument("--list", default="auto", choices=["auto", "half", "full"],
                    help="half = half/bin/atomonly/newton (the input scripts' default, newton on); full = newton off")
    ap.add_argument("--option", action="append", default=[], metavar="KEY=VALUE", help="b200md_set_option, e.g. halo=staged")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="skip secondary.* (development runs)")
    ap.add_argument("--no-parity", action="store_true", help="skip parity_check (development runs)")
    ap.add_argument("--no-e2e", action="store_true")
    a = ap.parse_args()
    ngpus = a.gpus
    cores = os.cpu_count() or 1
    kind, xyz, natoms, text = workload_desc(a.workload, ngpus)
    # `config` is the same object in both arms (the driver compares them); run-specific details go to `run`
    config = {"workload": text, "atoms_per_gpu": natoms // ngpus,
              "l2": "working set (neighbor list >= 0.5 GB/GPU) exceeds the 126 MB L2; no flush needed"}

    if a.impl == "reference":
        if int(os.environ.get("RANK", "0")) != 0:
            return 0
        val, lt, atoms = run_reference(kind, xyz, a.warmup, a.steps, cores)
        line = {"impl": "reference", "metric": "atom-steps/s", "value": val, "unit": "atom-steps/s", "n_gpus": ngpus,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": lt / a.steps * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": val, "unit": "atom-steps/s", "cores": cores, "kind": "reference",
                                 "sample": "full workload, %d steps, oracle/_ref/lmp -sf omp -pk omp %d" % (a.steps, cores)},
                "e2e": {"value": val, "unit": "atom-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        if not a.no_secondary:
            sec = {}
            for name, wl in (("eam", "eam"), ("eam_alloy_strong", "c5")):
                k2, xyz2, n2, text2 = workload_desc(wl, ngpus)
                sample = min(a.steps, 20)
                try:
                    v2, lt2, at2 = run_reference(k2, xyz2, 0, sample, cores)
                    sec[name] = {"value": v2, "unit": "atom-steps/s", "ms_per_step": lt2 / sample * 1e3, "workload": text2,
                                 "sample": "%d steps, oracle/_ref/lmp -sf omp -pk omp %d" % (sample, cores)}
                except Exception as e:
                    sec[name] = {"value": None, "unavailable": str(e)[:200]}
            line["secondary"] = sec
        print(json.dumps(line))
        return 0

    job = Job()
    B, rank, world = job.B, job.rank, job.world
    options = [tuple(o.split("=", 1)) for o in a.option]
    if a.lanes:
        options.append(("lanes", str(a.lanes)))
    run_info = {"decomposition": "x".join(str(g) for g in B.procgrid(world, [20.0 * xyz[0], 20.0 * xyz[1], 20.0 * xyz[2]]))}

    # ---- list kind.  The scripts run with `newton on` -> half list (f_j, rho_j by FP64 RED + reverse_comm), which is also
    # what measures fastest on B200 at every rank count (profiles/r02_summary.md); `--list full` = newton off ----
    lk_main = a.list if a.list != "auto" else "half"
    progress("start: %d rank(s), workload %s" % (world, a.workload))
    ctx = job.open(options)
    run_info["halo"] = ctx.get_option("halo")
    run_info["cpu_affinity"] = job.affinity
    ctx.close()
    run_info["neighbor_list"] = "half/bin/atomonly/newton" if lk_main == "half" else "full/bin/atomonly (newton off)"
    lk_sec = a.list if a.list != "auto" else "half"

    # ---- parity check on the N ranks, before anything is timed ----
    parity = {"nranks": world, "halo": run_info["halo"], "lists": {"lj": lk_main, "eam": lk_sec}}
    small = {}
    if not a.no_parity:
        for wl, lk in (("lj", lk_main), ("eam", lk_sec)):
            progress("parity check " + wl + " 32k, list " + lk)
            ctx = job.open(options)
            cfg = job.configure(ctx, wl, (1, 1, 1), lk)
            ctx.setup()
            recs = {0: ctx.thermo()}
            for t in ctx.run(100, cfg["dt"], 50):
                recs[t["step"]] = t
            bad = thermo_mismatch(recs, LOG[wl], 32000.0 if wl == "lj" else 1.0)
            parity[wl] = "ok" if bad is None else "MISMATCH " + bad
            # the 32 000-atom configs as shipped (C1, C2): 100 more steps, timed on the device
            job.barrier()
            ctx.run(100, cfg["dt"], 50 if wl == "eam" else 0)
            tm = ctx.timing()
            ms = job.maxr(tm["total_ms"])
            small[wl + "_32k"] = {"value": 32000 * 100 / (ms / 1e3), "unit": "atom-steps/s", "ms_per_step": ms / 100,
                                  "launches": int(tm["launches"]), "rebuilds": int(tm["nbuilds"]), "list": lk,
                                  "workload": "bench/in.%s as shipped (32 000 atoms)" % wl}
            ctx.close()
        # bench/in.chain (config C3: FENE bonds + lj/cut 1.12, special_bonds fene), 32 000 beads, without its thermostat
        # (fix langevin draws per rank in local atom order: one-rank only): 10 x `run 10` like the golden run of the
        # unmodified reference (tests/golden/chain.npz: T, PE, KE, Press, E_vdwl, E_mol), 1e-8 relative
        progress("parity check chain 32k")
        gold = np.load(os.path.join(ROOT, "tests", "golden", "chain.npz"))
        ctx = job.open(options)
        cfgc = job.P.inputs.in_chain(ctx, None, 1, 1, 1, list_kind=B.LIST_HALF_NEWTON if lk_main == "half" else B.LIST_FULL)
        ctx.setup()
        recs = [ctx.thermo()]
        for _ in range(10):
            ctx.setup()          # every `run` of the golden script starts with a set-up
            recs.append(ctx.run(10, cfgc["dt"], 0)[-1])
        worst = 0.0
        for t, w in zip(recs, [gold["thermo0"]] + list(gold["thermo_series"])):
            got = [t["temp"], t["pe"], t["ke"], t["press"], t["pe"] - t["emol"], t["emol"]]
            worst = max(worst, max(abs(x - y) / abs(y) for x, y in zip(got, w)))
        parity["chain"] = "ok" if worst <= 1e-8 else "MISMATCH worst relative deviation %.2e" % worst
        ctx.close()
        if any(str(parity[w]).startswith("MISMATCH") for w in ("lj", "eam", "chain")):
            if rank == 0:
                print(json.dumps({"parity_check": parity, "error": "thermo differs from the reference logs"}))
            return 2

    # ---- the workload reported as `value`: device-resident loop ----
    progress("main workload, list " + lk_main)
    ctx = job.open(options)
    cfg = job.configure(ctx, kind, xyz, lk_main)
    ctx.setup()
    m = measure(job, ctx, cfg, natoms, a.steps, a.warmup, ClockSampler(job.local_rank))
    nloc = ctx.nlocal()[0]
    roofline = roofline_of(kind, lk_main, m, nloc, ngpus, ctx.neigh_layout())

    # ---- end to end through the C ABI with HOST buffers (pinned), copies inside the timed region, every step.
    # "offload" (the headline e2e.value): the per-step boundary of the reference's own GPU package — host positions up,
    # forces down (src/GPU/pair_lj_cut_gpu.cpp:98-130).  The device integrates here, so the new positions come down too:
    # x up (24 B/atom); x, f down (48 B/atom).  b200md_get_atoms_async copies on a copy stream, x first and in pieces; the
    # host sends every piece of x back up (b200md_put_atoms_range) as soon as it has landed, so upload and download use the
    # two directions of the link at once, and the f download runs under the next step's kernels (host f buffers
    # alternate).  "offload_async": whole-array upload after the whole x download (round 2's first version);
    # "offload_sync": the same bytes, one blocking copy after the other.
    # "full_state": x, v up and x, v, f, tag down (a host that owns the whole state; what round 1 reported as e2e).
    e2e = None
    progress("main workload measured: %.4f ms/step; e2e" % m["ms_per_step"])
    if not a.no_e2e:
        n = ctx.nlocal()[0]
        cap = int(n * 1.05) + 1024
        hx, hv = B.pinned_empty((cap, 3)), B.pinned_empty((cap, 3))
        hf = [B.pinned_empty((cap, 3)), B.pinned_empty((cap, 3))]
        htag = B.pinned_empty((cap,), np.int32)
        ctx.get_atoms_into(hx, hv, hf[0], htag)
        variants = {}

        def step_full(it):
            ctx.put_atoms(hx, hv); ctx.run(1, cfg["dt"], 0); ctx.get_atoms_into(hx, hv, hf[0], htag)

        def step_sync(it):
            ctx.put_atoms(hx, None); ctx.run(1, cfg["dt"], 0); ctx.get_atoms_into(hx, None, hf[0], None)

        def step_async(it):
            ctx.put_atoms(hx, None)                 # H2D on the context's stream (next to the previous step's f download)
            ctx.run(1, cfg["dt"], 0)
            ctx.get_atoms_async(hx, hf[it & 1])     # pack, then D2H x, f on the copy stream
            ctx.wait_download(False)                # x is on the host: the next upload may start

        def step_stream(it):
            # the same bytes, streamed: every piece of x goes back up as soon as it has come down, so both directions of
            # the link are busy at once; the f download runs behind the x pieces and under the next step's kernels
            ctx.run(1, cfg["dt"], 0)
            ctx.get_atoms_async(hx, hf[it & 1])
            p = 0
            while True:
                first, count = ctx.wait_download_part(p)
                if count == 0:
                    break
                ctx.put_atoms_range(hx, first, count)
                p += 1

        for name, fn, up, down in (("full_state", step_full, 48, 76), ("offload_sync", step_sync, 24, 48),
                                   ("offload_async", step_async, 24, 48), ("offload", step_stream, 24, 48)):
            for it in range(3):
                fn(it)
            ctx.wait_download(True)
            job.barrier()
            t0 = time.perf_counter()
            h2d = d2h = 0
            for it in range(a.steps):
                n = ctx.nlocal()[0]
                fn(it)
                h2d += up * n; d2h += down * ctx.nlocal()[0]
            ctx.wait_download(True)
            job.barrier()
            ms = job.maxr((time.perf_counter() - t0) * 1e3)
            variants[name] = {"value": natoms * a.steps / (ms / 1e3), "ms_per_step": ms / a.steps,
                              "h2d_bytes_per_step": h2d // a.steps, "d2h_bytes_per_step": d2h // a.steps}
            ctx.get_atoms_into(hx, hv, hf[0], htag)
        e2e = dict(variants["offload"], unit="atom-steps/s",
                   api="every step: b200md_run(1) + b200md_get_atoms_async(x, f) + per piece of x: b200md_wait_download_part + "
                       "b200md_put_atoms_range; pinned host buffers, upload and download share the link, f download overlaps "
                       "the next step",
                   offload_async=dict(variants["offload_async"], api="b200md_put_atoms(x) + b200md_run(1) + b200md_get_atoms_async(x,f) "
                                      "+ b200md_wait_download(x): whole-array copies, f download overlapped"),
                   offload_sync=dict(variants["offload_sync"], api="b200md_put_atoms(x) + b200md_run(1) + b200md_get_atoms(x,f)"),
                   full_state=dict(variants["full_state"], api="b200md_put_atoms(x,v) + b200md_run(1) + b200md_get_atoms(x,v,f,tag)"))
    ctx.close()

    # ---- secondary workloads (same JSON line: the driver only ever runs `bench.py --gpus N`) ----
    secondary = dict(small)
    if not a.no_secondary:
        for name, wl in (("eam", "eam"), ("eam_alloy_strong", "c5")):
            if wl == a.workload:
                continue
            k2, xyz2, n2, text2 = workload_desc(wl, ngpus)
            progress("secondary " + name)
            ctx = job.open(options)
            cfg2 = job.configure(ctx, k2, xyz2, lk_sec)
            ctx.setup()
            entry = {"unit": "atom-steps/s", "workload": text2, "list": lk_sec,
                     "scaling": "strong" if wl == "c5" else "weak"}
            if wl == "c5" and not a.no_parity:
                # parity of the strong-scaling config from the run itself: 100 steps from the set-up, thermo 50
                gold = json.load(open(os.path.join(ROOT, "tests", "golden", "c5_thermo.json")))
                recs = {0: ctx.thermo()}
                for t in ctx.run(100, cfg2["dt"], 50):
                    recs[t["step"]] = t
                bad = None
                for step in ("0", "50", "100"):
                    for key in ("temp", "pe", "ke", "press"):
                        got, want = recs[int(step)][key], gold["thermo"][step][key]
                        if abs(got - want) > 1e-8 * abs(want) and bad is None:
                            bad = "step %s %s: got %.12g, reference %.12g" % (step, key, got, want)
                parity["eam_alloy_4M"] = "ok" if bad is None else "MISMATCH " + bad
                entry["rebuilds_first_100_steps"] = int(ctx.timing()["nbuilds"])
                entry["reference_rebuilds_first_100_steps"] = gold["neighbor_list_builds"]
            m2 = measure(job, ctx, cfg2, n2, a.steps, a.warmup)
            entry.update(value=m2["value"], ms_per_step=m2["ms_per_step"], launches=m2["launches"], rebuilds=m2["rebuilds"],
                         roofline=roofline_of(k2, lk_sec, m2, ctx.nlocal()[0], ngpus, ctx.neigh_layout()))
            ctx.close()
            secondary[name] = entry
        # config C3 scaled like in.chain.scaled to 2 048 000 beads per GPU; with `fix langevin` as shipped on one rank (the
        # reference's RanMars stream, sequence-exact), without it on several (the stream is per rank)
        progress("secondary chain")
        lang = world == 1
        ctx = job.open(options)
        cfgc = job.P.inputs.in_chain(ctx, None, *SCALE.get(ngpus, (4 * ngpus, 4, 4)), list_kind=B.LIST_HALF_NEWTON if lk_sec == "half" else B.LIST_FULL,
                                     langevin=lang)
        ctx.setup()
        nch = cfgc["natoms"]
        mc = measure(job, ctx, cfgc, nch, a.steps, a.warmup)
        secondary["chain"] = {"unit": "atom-steps/s", "value": mc["value"], "ms_per_step": mc["ms_per_step"], "launches": mc["launches"],
                              "rebuilds": mc["rebuilds"], "phases_ms_per_step": mc["phases"], "scaling": "weak", "list": lk_sec,
                              "workload": "bench/in.chain.scaled, %d beads (FENE + lj/cut 1.12, special_bonds fene), %s"
                                          % (nch, "fix langevin 1.0 1.0 10.0 904297 as shipped" if lang else "NVE (fix langevin left out on several ranks)")}
        ctx.close()
        if str(parity.get("eam_alloy_4M", "ok")).startswith("MISMATCH"):
            if rank == 0:
                print(json.dumps({"parity_check": parity, "error": "eam/alloy 4M thermo differs from the reference"}))
            return 2

    line = {"metric": "atom-steps/s", "value": m["value"], "unit": "atom-steps/s", "n_gpus": ngpus, "steps": a.steps,
            "warmup": max(a.warmup, 3), "ms_per_step": m["ms_per_step"], "higher_is_better": True,
            "scaling": "strong" if a.workload == "c5" else "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config, "run": run_info,
            "parity_check": parity, "roofline": roofline, "e2e": e2e, "secondary": secondary,
            "gpu_launches": m["launches"], "clocks": m["clocks"], "device_ms": m["device_ms"], "wall_ms": m["wall_ms"],
            "rebuilds_in_timed_region": m["rebuilds"]}
    progress("done with the GPU part")
    if rank == 0 and not a.no_cpu_baseline and ngpus == 1:
        try:
            os.sched_setaffinity(0, range(os.cpu_count() or 1))    # the CPU reference gets every host core again
            sample_steps = 20
            val, lt, atoms = run_reference(kind, xyz, 0, sample_steps, cores)
            line["cpu_baseline"] = {"value": val, "unit": "atom-steps/s", "cores": cores, "kind": "reference",
                                    "sample": "same workload (%d atoms), %d steps, oracle/_ref/lmp -sf omp -pk omp %d, Loop time %.3f s"
                                              % (atoms, sample_steps, cores, lt)}
        except Exception as e:  # the checker is optional for the measurement itself
            line["cpu_baseline"] = {"value": None, "unit": "atom-steps/s", "cores": cores, "kind": "reference",
                                    "sample": "unavailable: %s" % str(e)[:200]}
    if rank == 0:
        print(json.dumps(line))
    if job.dist is not None:
        job.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
