"""bench/in.chain on one GPU: thermo + timing.  python scripts/chain_run.py [steps] [half|full|<replicate>] [langevin]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200pkg
P = b200pkg.load(); B = P.b200md
steps = int(sys.argv[1]) if len(sys.argv) > 1 else 100
lk = B.LIST_FULL if (len(sys.argv) > 2 and sys.argv[2] == "full") else B.LIST_HALF_NEWTON
rep = int(sys.argv[2]) if (len(sys.argv) > 2 and sys.argv[2].isdigit()) else 1
ctx = B.Context(0)
cfg = P.inputs.in_chain(ctx, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "data", "data.chain"), rep, rep, rep, list_kind=lk, langevin="langevin" in sys.argv)
ctx.setup()
n = cfg["natoms"]
t = ctx.thermo()
print("step 0 T %.8g E_pair %.8g E_mol %.8g TotEng %.8g Press %.8g" % (t["temp"], (t["pe"] - t["emol"]) / n, t["emol"] / n, (t["pe"] + t["ke"]) / n, t["press"]))
ctx.run(20, cfg["dt"], 0)
t = ctx.run(steps, cfg["dt"], 0)[-1]
tm = ctx.timing(); st = ctx.neigh_stats()
print("step %d T %.8g E_pair %.8g E_mol %.8g TotEng %.8g Press %.8g" % (t["step"], t["temp"], (t["pe"] - t["emol"]) / n, t["emol"] / n, (t["pe"] + t["ke"]) / n, t["press"]))
print("chain %d atoms: %.4f ms/step %.1f Matom-steps/s builds %d launches %d neigh/atom %.2f" % (n, tm["total_ms"] / steps, n * steps / tm["total_ms"] / 1e3, tm["nbuilds"], tm["launches"], st["total"] / n))
ctx.close()
