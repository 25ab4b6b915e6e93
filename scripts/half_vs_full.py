"""Full list / newton off (device default) against half list / newton on (the reference's CPU
default; FP64 RED for f_j and rho_j) at 2 048 000 atoms: 100 steps each."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200pkg
P = b200pkg.load(); B = P.b200md

def run(which, kind, tpa, x=4, steps=100):
    ctx = B.Context(0)
    ctx.set_lanes_per_atom(tpa)
    if which == "lj":
        cfg = P.inputs.in_lj(ctx, x, x, x, list_kind=kind)
    else:
        cfg = P.inputs.in_eam(ctx, x, x, x, potential="oracle/_ref/potentials/Cu_u3.eam", list_kind=kind)
    ctx.setup()
    t0 = ctx.thermo()
    th = ctx.run(steps, cfg["dt"], 0)[-1]
    tm = ctx.timing()
    print("%s %s tpa=%d: %.3f ms/step %.1f Matom-steps/s | step0 pe %.10g press %.10g | step%d T %.8g pe %.10g press %.8g" % (
        which, "half/newton" if kind == B.LIST_HALF_NEWTON else "full", tpa, tm["total_ms"] / steps,
        cfg["natoms"] * steps / tm["total_ms"] / 1e3, t0["pe"], t0["press"], steps, th["temp"], th["pe"], th["press"]), flush=True)
    ctx.close()

for which in ("lj", "eam"):
    for kind in (B.LIST_FULL, B.LIST_HALF_NEWTON):
        for tpa in (1, 2, 4, 8):
            run(which, kind, tpa)
