#!/usr/bin/env python3
"""bench.py — atom-steps/s of the short-range MD hot path (neighbor build -> lj/cut | eam -> fix nve)
on N B200s of one node, measured on the reference's own bench inputs.

  python bench.py --gpus N --steps K --warmup W [--workload lj|eam] [--impl reference]

A "step" is one MD timestep of the workload (Verlet::run body, /root/reference/src/verlet.cpp:246-359)
including the neighbor rebuilds that fall inside the K steps.  Workload at N GPUs: bench/in.lj scaled to
2 048 000 atoms per GPU (-var x/y/z = 4,4,4 / 8,4,4 / 8,8,4 / 8,8,8), weak scaling, one process per GPU,
spatial decomposition with NCCL halo exchange.  One JSON line on rank 0.
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
# algorithmic bytes: force pass per atom per launch, and whole step per atom-step
N_FULL = {"lj": 75.2, "eam": 75.5, "chain": 9.74}
BYTES_KERNEL = {"lj": 28 * 75.2 + 52, "eam": 64 * 75.5 + 104,    # LJ force pass; EAM rho+fp and force passes
                "chain": 28 * 9.74 + 52 + 1 * (8 + 24 + 24) + 52}  # + FENE: 0.99 bonds/atom x (tag,map + x_j + f_j RED) + own x,f
# in.chain (SURVEY's method applied to C3, N=9.74, rebuild every ~4 steps): force passes 381 + NVE 192
# + check_distance 48 + rebuild (4N list + 24 xhold + 176 sort + 28N candidates + 8 map)/4 = 128
# fix langevin adds 12 B written + 12 B read of random words, 4 B tag, 4 B rank per atom-step (its v/f
# streams are the ones nve_final moves anyway: the two are one fused pass)
BYTES_STEP = {"lj": 2480.0, "eam": 5510.0, "chain": 750.0, "chain+langevin": 782.0}
# (the figures above are SURVEY 8(d)'s yardstick, defined on the full list; the half-list kernels do
#  the same physical work -- every pair once instead of twice -- and are rated with the same figure)
# measured choice on ONE GPU (profiles/r01_summary.md): the half/newton list wins everywhere.  On several
# GPUs the better list kind depends on how much of the halo the force kernel hides (full: no reverse pass;
# half: half the pair work), so `--list auto` MEASURES it there: 20 untimed steps of each kind before the
# warm-up, the faster one (max over ranks) runs the timed region.  The table's N>1 column is only the
# order in which the candidates are tried.
LIST_AUTO = {"lj": {1: "half", 2: "full"}, "eam": {1: "half", 2: "half"}, "chain": {1: "half", 2: "half"}}
# algorithmic bytes of the half-list kernels per atom per step (N_h listed pairs per atom, each:
# 4 index + 24 x_j [+ 8 fp_j] gathered, 24 f_j [8 rho_j] as RED payload; plus the per-atom terms)
N_HALF = {"lj": 37.6, "eam": 37.75, "chain": 4.87}
BYTES_KERNEL_HALF = {"lj": 52 * 37.6 + 52, "eam": (36 + 60) * 37.75 + 104, "chain": 52 * 4.87 + 52 + 56 + 52}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


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


def ref_script(workload, xyz, warm, steps, langevin=True):
    from oracle import oracle as O
    if workload == "lj":
        body = O.IN_LJ.format(nx=20 * xyz[0], ny=20 * xyz[1], nz=20 * xyz[2])
    elif workload == "chain":   # bench/in.chain.scaled (what the B200 arm runs, thermostat included unless --no-langevin)
        body = ("units lj\natom_style bond\nspecial_bonds fene\nread_data %s\nreplicate %d %d %d\n"
                "neighbor 0.4 bin\nneigh_modify every 1 delay 1\nbond_style fene\nbond_coeff 1 30.0 1.5 1.0 1.0\n"
                "pair_style lj/cut 1.12\npair_modify shift yes\npair_coeff 1 1 1.0 1.0 1.12\nfix 1 all nve\n"
                "%stimestep 0.012\n" % (os.path.join(os.path.dirname(O.POT_DIR), "data", "data.chain"), *xyz,
                                       "fix 2 all langevin 1.0 1.0 10.0 904297\n" if langevin else ""))
    else:
        body = O.IN_EAM.format(nx=20 * xyz[0], ny=20 * xyz[1], nz=20 * xyz[2], style="eam",
                               coeff="1 1 " + os.path.join(O.POT_DIR, "Cu_u3.eam"))
    body = body.replace("atom_modify map array\n", "")
    return body + ("run %d\n" % warm if warm > 0 else "") + "run %d\n" % steps


def run_reference(workload, xyz, warm, steps, cores, langevin=True):
    """the reference's own CPU implementation (oracle/_ref/lmp, OPENMP package on all host cores)"""
    import re
    from oracle import oracle as O
    t0 = time.time()
    _, atoms, nst, out = O.ref_exe_run(ref_script(workload, xyz, warm, steps, langevin), threads=cores, omp=True, timeout=3000)
    loops = re.findall(r"Loop time of ([0-9.eE+-]+) on \d+ procs for (\d+) steps with (\d+) atoms", out)
    lt, nst, atoms = float(loops[-1][0]), int(loops[-1][1]), int(loops[-1][2])
    return atoms * nst / lt, lt, atoms, time.time() - t0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workload", default="lj", choices=["lj", "eam", "chain"])
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--lanes", type=int, default=0)
    ap.add_argument("--list", default="auto", choices=["auto", "half", "full"],
                    help="half = half/bin/atomonly/newton (the input scripts' default, newton on); full = newton off")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-langevin", action="store_true",
                    help="chain workload: leave `fix langevin` out (one GPU runs it by default, as bench/in.chain ships it)")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    ngpus = a.gpus
    xyz = SCALE.get(ngpus, (4 * ngpus, 4, 4))
    natoms = 4 * 8000 * xyz[0] * xyz[1] * xyz[2]
    wl = a.workload
    # the thermostat's random stream follows one rank's local atom order (the reference seeds every rank
    # separately): it is part of the chain workload on one GPU
    langevin = wl == "chain" and not a.no_langevin and world == 1 and ngpus == 1
    desc = {"lj": "fcc lattice, lj/cut 2.5, neigh every 20", "eam": "fcc lattice, eam Cu_u3, neigh check yes",
            "chain": "data.chain replicated, FENE bonds + lj/cut 1.12, special_bonds fene, "
                     + ("fix langevin 1.0 1.0 10.0 904297 (the reference's RanMars stream)" if langevin else "fix langevin left out")}[wl]
    config = {"workload": "bench/in.%s -var x %d -var y %d -var z %d (%d atoms, %s)" % (
        wl + (".scaled" if wl == "chain" else ""), xyz[0], xyz[1], xyz[2], natoms, desc + (", NVE" if not langevin else "")),
        "atoms_per_gpu": natoms // ngpus, "decomposition": None,
        "l2": "working set (neighbor list >= 1 GB/GPU) exceeds the 126 MB L2; no flush needed"}
    cores = os.cpu_count() or 1

    if a.impl == "reference":
        if rank != 0:
            return 0
        val, lt, atoms, wall = run_reference(wl, xyz, a.warmup, a.steps, cores, langevin)
        line = {"impl": "reference", "metric": "atom-steps/s", "value": val, "unit": "atom-steps/s", "n_gpus": ngpus,
                "steps": a.steps, "warmup": a.warmup, "ms_per_step": lt / a.steps * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": val, "unit": "atom-steps/s", "cores": cores, "kind": "reference",
                                 "sample": "full workload, %d steps, oracle/_ref/lmp -sf omp -pk omp %d" % (a.steps, cores)},
                "e2e": {"value": val, "unit": "atom-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    import b200pkg
    P = b200pkg.load()
    B = P.b200md
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    ctx = B.Context(local_rank, rank, world)
    if world > 1:
        import torch
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.frombuffer(bytearray(B.Context.unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(idt, 0)
        ctx.comm_init(bytes(idt.cpu().numpy().tobytes()))
    if a.lanes:
        ctx.set_lanes_per_atom(a.lanes)
    # list kind: the scripts run with `newton on` -> half list (f_j, rho_j by FP64 RED + reverse_comm);
    # measured on B200 (profiles/): half wins on one GPU; with several ranks the full list lets the
    # force kernel overlap the halo and needs no reverse_comm
    lk = a.list if a.list != "auto" else LIST_AUTO[wl][min(world, 2)]
    list_kind = B.LIST_HALF_NEWTON if lk == "half" else B.LIST_FULL
    config["neighbor_list"] = "half/bin/atomonly/newton" if lk == "half" else "full/bin/atomonly (newton off)"
    if wl == "lj":
        cfg = P.inputs.in_lj(ctx, *xyz, list_kind=list_kind)
    elif wl == "chain":
        cfg = P.inputs.in_chain(ctx, os.path.join(ROOT, "oracle", "_ref", "data", "data.chain"), *xyz, list_kind=list_kind,
                                langevin=langevin)
    else:
        cfg = P.inputs.in_eam(ctx, *xyz, potential=os.path.join(ROOT, "oracle", "_ref", "potentials", "Cu_u3.eam"),
                              list_kind=list_kind)
    config["decomposition"] = "x".join(str(g) for g in B.procgrid(world, [20.0 * xyz[0], 20.0 * xyz[1], 20.0 * xyz[2]]))
    config["halo"] = ("none (one rank)" if world == 1 else
                      "staged (3 x 2 NCCL send/recv stages)" if os.environ.get("B200MD_HALO") == "staged" else
                      "direct (owner <-> holder, one NCCL group per exchange)")
    ctx.setup()

    def barrier():
        if dist is not None:
            import torch
            dist.barrier()
            torch.cuda.synchronize()

    def maxr(v):
        if dist is None:
            return v
        import torch
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- several ranks, --list auto: the list kind is chosen by measurement (untimed, before the warm-up) ----
    if world > 1 and a.list == "auto" and wl != "chain":
        trial = {}
        for kind in ("full", "half"):
            ctx.neigh_config(**cfg["neigh"], list_kind=B.LIST_HALF_NEWTON if kind == "half" else B.LIST_FULL)
            ctx.setup()
            ctx.run(5, cfg["dt"], 0)
            barrier()
            ctx.run(20, cfg["dt"], 0)
            trial[kind] = maxr(ctx.timing()["total_ms"]) / 20     # CUDA events on the context's stream, max over ranks
        lk = min(trial, key=trial.get)
        list_kind = B.LIST_HALF_NEWTON if lk == "half" else B.LIST_FULL
        ctx.neigh_config(**cfg["neigh"], list_kind=list_kind)
        ctx.setup()
        config["neighbor_list"] = "half/bin/atomonly/newton" if lk == "half" else "full/bin/atomonly (newton off)"
        config["list_trial_ms_per_step"] = trial

    # ---- device-resident loop: W warm-up steps, K timed steps ----
    ctx.run(max(a.warmup, 3), cfg["dt"], 0)
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    t0 = time.perf_counter()
    ctx.run(a.steps, cfg["dt"], 0)
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    tm = ctx.timing()
    clocks = sampler.result()
    dev_ms = maxr(tm["total_ms"])
    ms = dev_ms   # device time (CUDA events around the K steps on the context's stream), max over ranks; wall_ms is reported beside it
    value = natoms * a.steps / (ms / 1e3)

    # ---- dominant kernel: pair force kernel(s), CUDA events around each launch on its stream ----
    os.environ["B200MD_PHASE_TIMING"] = "1"
    ctx.run(20, cfg["dt"], 0)
    os.environ["B200MD_PHASE_TIMING"] = "0"
    tp = ctx.timing()
    n_eval = 20
    pair_ms = tp["pair_ms"] / n_eval  # per timestep (LJ: one launch; EAM: rho+fp, fp halo, force)
    nloc = ctx.nlocal()[0]
    hbm, which = peaks()
    bytes_kernel = BYTES_KERNEL_HALF[wl] if lk == "half" else BYTES_KERNEL[wl]
    bstep = BYTES_STEP["chain+langevin" if langevin else wl]
    achieved = bytes_kernel * nloc / (pair_ms / 1e3) / 1e9
    roofline = {"bound": "hbm", "kernel": ("k_lj" if wl == "lj" else "k_lj+k_bond_fene" if wl == "chain" else ("k_eam_rho_half+k_eam_embed+k_eam_force" if lk == "half" else "k_eam_rho_fp+k_eam_force")),
                "list": lk, "achieved": achieved,
                "peak": hbm, "peak_source": which, "unit": "GB/s", "frac": achieved / hbm, "traffic": None,
                "kernel_ms": pair_ms, "bytes_per_atom_launch": bytes_kernel,
                "step_achieved": bstep * (value / ngpus) / 1e9, "step_frac": bstep * (value / ngpus) / 1e9 / hbm,
                "phases_ms_per_step": {k: tp[k + "_ms"] / n_eval for k in ("pair", "neigh", "comm", "modify")}}
    tr = os.path.join(ROOT, "profiles", "traffic_%s_%s.json" % (wl, lk))
    if not os.path.exists(tr):
        tr = os.path.join(ROOT, "profiles", "traffic_%s.json" % wl)
    if os.path.exists(tr):
        roofline["traffic"] = json.load(open(tr)).get("dram_bytes_per_launch")

    # ---- end to end through the C ABI with HOST buffers: every step uploads x,v (pinned) and
    # downloads x,v,f,tag of the step's result ----
    n = ctx.nlocal()[0]
    cap = int(n * 1.05) + 1024
    hx, hv, hf = B.pinned_empty((cap, 3)), B.pinned_empty((cap, 3)), B.pinned_empty((cap, 3))
    htag = B.pinned_empty((cap,), np.int32)
    ctx.get_atoms_into(hx, hv, hf, htag)
    e2e_steps = a.steps
    for _ in range(3):
        ctx.put_atoms(hx, hv); ctx.run(1, cfg["dt"], 0); ctx.get_atoms_into(hx, hv, hf, htag)
    barrier()
    t0 = time.perf_counter()
    h2d = d2h = 0
    for _ in range(e2e_steps):
        n = ctx.nlocal()[0]
        ctx.put_atoms(hx, hv)
        ctx.run(1, cfg["dt"], 0)
        n = ctx.nlocal()[0]
        ctx.get_atoms_into(hx, hv, hf, htag)
        h2d += 48 * n; d2h += 76 * n
    barrier()
    e2e_ms = maxr((time.perf_counter() - t0) * 1e3)
    e2e = {"value": natoms * e2e_steps / (e2e_ms / 1e3), "unit": "atom-steps/s", "h2d_bytes_per_step": h2d // e2e_steps,
           "d2h_bytes_per_step": d2h // e2e_steps, "ms_per_step": e2e_ms / e2e_steps,
           "api": "b200md_put_atoms(x,v) + b200md_run(1) + b200md_get_atoms(x,v,f,tag), pinned host buffers"}

    line = {"metric": "atom-steps/s", "value": value, "unit": "atom-steps/s", "n_gpus": ngpus, "steps": a.steps,
            "warmup": max(a.warmup, 3), "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config, "roofline": roofline, "e2e": e2e,
            "gpu_launches": int(tm["launches"]), "clocks": clocks, "device_ms": dev_ms, "wall_ms": maxr(wall_ms),
            "rebuilds_in_timed_region": int(tm["nbuilds"])}
    if rank == 0 and not a.no_cpu_baseline and ngpus == 1:
        try:
            sample_steps = 20
            val, lt, atoms, wall = run_reference(wl, xyz, 0, sample_steps, cores, langevin)
            line["cpu_baseline"] = {"value": val, "unit": "atom-steps/s", "cores": cores, "kind": "reference",
                                    "sample": "same workload (%d atoms), %d steps, oracle/_ref/lmp -sf omp -pk omp %d, Loop time %.3f s"
                                              % (atoms, sample_steps, cores, lt)}
        except Exception as e:  # the checker is optional for the measurement itself
            line["cpu_baseline"] = {"value": None, "unit": "atom-steps/s", "cores": cores, "kind": "reference",
                                    "sample": "unavailable: %s" % str(e)[:200]}
    ctx.close()
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
