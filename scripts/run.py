"""The one development runner (everything else the round needs is bench.py and pytest).

  python scripts/run.py lj|eam|c5|chain [--steps K] [--x X] [--list half|full] [--lanes L] [--langevin]
                                        [--option key=value ...] [--phases]
      one workload on one GPU (or under torchrun on several): thermo of the last step + timing; short enough to
      sit under ncu (`ncu ... python scripts/run.py lj --steps 3`)

  bash-level recipes (B200_PROFILING.md), with WL in lj|eam|c5|chain:
      launch list : python scripts/run.py WL --steps 25 && ncu --metrics gpu__time_duration.sum --clock-control none
                    -c 600 --csv --log-file gpurun_out/launches_WL.csv python scripts/run.py WL --steps 25
      full capture: python scripts/run.py WL --steps 3 && ncu --set full --clock-control none --import-source on
                    -k regex:'k_lj|k_build' -s 1 -c 3 -o gpurun_out/prof_WL python scripts/run.py WL --steps 3
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200pkg  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("workload", choices=["lj", "eam", "c5", "chain"])
ap.add_argument("--steps", type=int, default=25)
ap.add_argument("--x", type=int, default=0, help="-var x/y/z (default: 4 for lj/eam/chain, 5 for c5)")
ap.add_argument("--xyz", default="", help="-var x,y,z separately, e.g. 8,4,4 (weak scaling on 2 ranks)")
ap.add_argument("--list", default="half", choices=["half", "full"])
ap.add_argument("--lanes", type=int, default=0)
ap.add_argument("--langevin", action="store_true")
ap.add_argument("--phases", action="store_true")
ap.add_argument("--trace", type=int, default=0, help="after the run: N more steps with both streams' milestones printed (option trace)")
ap.add_argument("--option", action="append", default=[])
a = ap.parse_args()
P = b200pkg.load(); B = P.b200md
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dist = None
if world > 1:   # torch (and its bundled NCCL) is loaded before the library binds libnccl.so.2
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
ctx = B.Context(lr, rank, world)
if world > 1:
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt = torch.frombuffer(bytearray(B.Context.unique_id()), dtype=torch.uint8).cuda()
    dist.broadcast(idt, 0)
    ctx.comm_init(bytes(idt.cpu().numpy().tobytes()))
for o in a.option:
    ctx.set_option(*o.split("=", 1))
if a.lanes:
    ctx.set_lanes_per_atom(a.lanes)
if a.phases:
    ctx.set_option("phase_timing", "on")
lk = B.LIST_FULL if a.list == "full" else B.LIST_HALF_NEWTON
x = a.x or (5 if a.workload == "c5" else 4)
x, y, z = [int(v) for v in a.xyz.split(",")] if a.xyz else (x, x, x)
if a.workload == "lj":
    cfg = P.inputs.in_lj(ctx, x, y, z, list_kind=lk)
elif a.workload == "eam":
    cfg = P.inputs.in_eam(ctx, x, y, z, list_kind=lk)
elif a.workload == "c5":
    cfg = P.inputs.in_eam(ctx, x, y, z, style="eam/alloy", list_kind=lk)
else:
    cfg = P.inputs.in_chain(ctx, None, x, y, z, list_kind=lk, langevin=a.langevin)
ctx.setup()
t0 = ctx.thermo()
th = ctx.run(a.steps, cfg["dt"], 0)[-1]
tm = ctx.timing()
ms = tm["total_ms"]
if dist is not None:
    t = torch.tensor([ms], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t)
if rank == 0:
    n = cfg["natoms"]
    if a.phases:
        print("phases ms/step: pair %.4f neigh %.4f comm %.4f modify %.4f" % tuple(tm[k] / a.steps for k in ("pair_ms", "neigh_ms", "comm_ms", "modify_ms")))
    print("%s %s list, %d atoms on %d GPU(s): %.4f ms/step, %.1f Matom-steps/s, %d rebuilds, %d launches | step 0 T %.8g PE %.10g P %.8g | step %d T %.8g PE %.10g P %.8g"
          % (a.workload, a.list, n, world, ms / a.steps, n * a.steps / ms / 1e3, tm["nbuilds"], tm["launches"], t0["temp"], t0["pe"], t0["press"],
             th["step"], th["temp"], th["pe"], th["press"]), flush=True)
if a.trace:
    ctx.set_option("trace", "on")
    ctx.run(a.trace + 1, cfg["dt"], 0)
ctx.close()
if dist is not None:
    dist.destroy_process_group()
