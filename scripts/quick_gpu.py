"""First-contact GPU check: bench/in.lj and bench/in.eam at 32k atoms against the thermo the
reference logs print (bench/log.6Oct16.{lj,eam}.fixed.icc.1), then timing sweeps."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import b200pkg
P = b200pkg.load(); B = P.b200md

def show(tag, th, per_atom):
    n = th["natoms"] if per_atom else 1.0
    print("%s step %d T %.8g E_pair %.8g TotEng %.8g Press %.8g" % (tag, th["step"], th["temp"], th["pe"] / n, (th["pe"] + th["ke"]) / n, th["press"]), flush=True)

def run(which, x, steps, tpa, pot=None, style="eam"):
    ctx = B.Context(0)
    ctx.set_lanes_per_atom(tpa)
    cfg = P.inputs.in_lj(ctx, x, x, x) if which == "lj" else P.inputs.in_eam(ctx, x, x, x, potential=pot, style=style)
    t0 = time.time(); ctx.setup(); t1 = time.time()
    show(which, ctx.thermo(), which == "lj")
    th = ctx.run(steps, cfg["dt"], cfg["thermo"])
    for t in th: show(which, t, which == "lj")
    tm = ctx.timing(); st = ctx.neigh_stats()
    print("%s n=%d tpa=%d setup %.3fs loop %.3f ms -> %.1f Matom-step/s builds %d neigh/atom %.2f pitch %d launches %d" % (
        which, cfg["natoms"], tpa, t1 - t0, tm["total_ms"], cfg["natoms"] * steps / tm["total_ms"] / 1e3, tm["nbuilds"],
        st["total"] / cfg["natoms"], st["pitch"], tm["launches"]), flush=True)
    if os.environ.get("B200MD_PHASE_TIMING") == "1":
        print("   phases ms: pair %.2f neigh %.2f comm %.2f modify %.2f" % (tm["pair_ms"], tm["neigh_ms"], tm["comm_ms"], tm["modify_ms"]))
    ctx.close()

pot = "oracle/_ref/potentials/Cu_u3.eam"
print("golden lj : step 0 T 1.44 E_pair -6.7733681 TotEng -4.6134356 P -5.0197073 | step 100 T 0.7574531 E_pair -5.7585055 TotEng -4.6223613 P 0.20726105")
run("lj", 1, 100, 8)
print("golden eam: step 0 T 1600 E_pair -113280 TotEng -106662.09 P 18703.573 | 50: 781.69049 -109873.35 -106640.13 52273.088 | 100: 801.832 -109957.3 -106640.77 51322.821")
run("eam", 1, 100, 8, pot)
if len(sys.argv) > 1 and sys.argv[1] == "sweep":
    for tpa in (4, 8, 16, 32):
        run("lj", 4, 100, tpa)
    for tpa in (4, 8, 16):
        run("eam", 4, 100, tpa, pot)
    run("eam", 2, 100, 8, "oracle/_ref/potentials/Cu_mishin1.eam.alloy", "eam/alloy")
