"""Short run for ncu: python scripts/profile_run.py lj|eam|eamalloy [steps] [lanes] [x] [half|full]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200pkg
P = b200pkg.load(); B = P.b200md
wl = sys.argv[1] if len(sys.argv) > 1 else "lj"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
lanes = int(sys.argv[3]) if len(sys.argv) > 3 else 0
x = int(sys.argv[4]) if len(sys.argv) > 4 else 4
lk = B.LIST_FULL if (len(sys.argv) > 5 and sys.argv[5] == "full") else B.LIST_HALF_NEWTON
ctx = B.Context(0)
if lanes: ctx.set_lanes_per_atom(lanes)
pot = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "potentials")
if wl == "lj": cfg = P.inputs.in_lj(ctx, x, x, x, list_kind=lk)
elif wl == "eam": cfg = P.inputs.in_eam(ctx, x, x, x, potential=os.path.join(pot, "Cu_u3.eam"), list_kind=lk)
else: cfg = P.inputs.in_eam(ctx, x, x, x, potential=os.path.join(pot, "Cu_mishin1.eam.alloy"), style="eam/alloy", list_kind=lk)
ctx.setup()
th = ctx.run(steps, cfg["dt"], 0)
tm = ctx.timing()
if os.environ.get("B200MD_PHASE_TIMING") == "1":
    print("phases ms/step: pair %.4f neigh %.4f comm %.4f modify %.4f" % tuple(tm[k] / steps for k in ("pair_ms", "neigh_ms", "comm_ms", "modify_ms")), end=" | ")
print(wl, cfg["natoms"], "steps", steps, "ms/step %.4f" % (tm["total_ms"] / steps), "Matom-steps/s %.1f" % (cfg["natoms"] * steps / tm["total_ms"] / 1e3), th[-1]["temp"])
ctx.close()
