"""torchrun -n N scripts/strong_eam.py [x] : config C5 — in.eam with pair_style eam/alloy Cu_mishin1,
-var x/y/z 5 (4 000 000 atoms), strong scaling over N GPUs with atom migration at every rebuild."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200pkg
P = b200pkg.load(); B = P.b200md
rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
x = int(sys.argv[1]) if len(sys.argv) > 1 else 5
if world > 1:  # torch (and its bundled NCCL) must be loaded before libb200md.so binds libnccl.so.2
    import torch, torch.distributed as dist
    torch.cuda.set_device(lr)
ctx = B.Context(lr, rank, world)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0: idt = torch.frombuffer(bytearray(B.Context.unique_id()), dtype=torch.uint8).cuda()
    dist.broadcast(idt, 0)
    ctx.comm_init(bytes(idt.cpu().numpy().tobytes()))
lk = B.LIST_HALF_NEWTON if (len(sys.argv) > 2 and sys.argv[2] == "half") else B.LIST_FULL
cfg = P.inputs.in_eam(ctx, x, x, x, style="eam/alloy", list_kind=lk)
ctx.setup()
t0 = ctx.thermo()
ctx.run(10, cfg["dt"], 0)
th = ctx.run(100, cfg["dt"], 50)
tm = ctx.timing(); nl, ng = ctx.nlocal()
if world > 1:
    t = torch.tensor([tm["total_ms"]], dtype=torch.float64, device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t)
else: ms = tm["total_ms"]
if rank == 0:
    print("eam/alloy Cu_mishin1 " + ("half" if lk == B.LIST_HALF_NEWTON else "full") + " %d atoms N=%d: %.3f ms/step %.1f Matom-steps/s builds %d rank0 nlocal %d nghost %d | step0 T %.8g PE %.10g P %.8g | step110 T %.8g PE %.10g P %.8g" % (
        cfg["natoms"], world, ms / 100, cfg["natoms"] * 100 / ms / 1e3, tm["nbuilds"], nl, ng, t0["temp"], t0["pe"], t0["press"], th[-1]["temp"], th[-1]["pe"], th[-1]["press"]), flush=True)
ctx.close()
if world > 1: dist.destroy_process_group()
