"""torchrun -n N scripts/mgpu_parity.py : bench/in.lj and in.eam (32 000 atoms) decomposed over N GPUs;
thermo must reproduce the reference logs (8 digits) exactly like the 1-GPU run."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import b200pkg
P = b200pkg.load(); B = P.b200md
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
def mk():
    ctx = B.Context(lr, rank, world)
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0: idt = torch.frombuffer(bytearray(B.Context.unique_id()), dtype=torch.uint8).cuda()
    dist.broadcast(idt, 0)
    ctx.comm_init(bytes(idt.cpu().numpy().tobytes()))
    return ctx
x = int(sys.argv[1]) if len(sys.argv) > 1 else 1
lk = B.LIST_HALF_NEWTON if (len(sys.argv) > 2 and sys.argv[2] == "half") else B.LIST_FULL
ok = True
for wl in ("lj", "eam"):
    ctx = mk()
    if wl == "lj": cfg = P.inputs.in_lj(ctx, x, x, x, list_kind=lk); per = cfg["natoms"]
    else: cfg = P.inputs.in_eam(ctx, x, x, x, list_kind=lk); per = 1.0
    ctx.setup()
    recs = [ctx.thermo()] + ctx.run(100, cfg["dt"], 50)
    nl, ng = ctx.nlocal()
    tm = ctx.timing()
    if rank == 0:
        for t in recs:
            print("%s N=%d step %d T %.8g E_pair %.8g TotEng %.8g Press %.8g" % (wl, world, t["step"], t["temp"], t["pe"] / per, (t["pe"] + t["ke"]) / per, t["press"]), flush=True)
        print("   rank0 nlocal %d nghost %d loop %.2f ms builds %d" % (nl, ng, tm["total_ms"], tm["nbuilds"]), flush=True)
    ctx.close()
# bench/in.chain without fix langevin (its stream is per rank): FENE bonds across sub-box faces, bond
# shares of ghost partners returned by reverse_comm; must equal tests/golden/chain.npz (the 1-rank reference)
if "chain" in sys.argv:
    import numpy as np
    g = np.load("tests/golden/chain.npz")
    ctx = mk()
    cfg = P.inputs.in_chain(ctx, list_kind=lk)
    ctx.setup()
    recs = [ctx.thermo()]
    for _ in range(10):       # every `run 10` of the golden run starts with a set-up (tests/test_chain.py)
        ctx.setup()
        recs.append(ctx.run(10, cfg["dt"], 0)[-1])
    if rank == 0:
        want = [g["thermo0"]] + list(g["thermo_series"])
        worst = 0.0
        for t, w in zip(recs, want):
            got = [t["temp"], t["pe"], t["ke"], t["press"], t["pe"] - t["emol"], t["emol"]]
            worst = max(worst, max(abs(a - b) / abs(b) for a, b in zip(got, w)))
        t = recs[-1]
        print("chain N=%d step %d T %.8g PE %.10g Press %.8g | worst relative deviation from the reference over 100 steps %.2e"
              % (world, t["step"], t["temp"], t["pe"], t["press"], worst), flush=True)
    ctx.close()
dist.destroy_process_group()
