"""A/B of differently built libraries on the all-pairs force pass (one GPU): for every library given, in a fresh
process (PSIM_LIBRARY selects it), N particles of the synthetic lattice, device-timed steps, and the filter audit
(candidates evaluated == ordered in-range pairs).

    python tools/ab_allpairs.py [--n 262144] [--precision 8] [--steps 20] [--repeat 2] name=path.so [name=path.so ...]
"""
import argparse
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child(args):
    sys.path.insert(0, ROOT)
    import particle_simulation_b200 as P
    from particle_simulation_b200 import generate as G
    ids, pos, vel, area = G.synthetic(args.n)
    sim = P.Simulation(P.make_params(length_of_area=area), ids, pos, vel, args.precision)
    sim.set_option(P.OPT_ALLPAIRS_KERNEL, args.variant)
    if args.cells:
        sim.set_option(P.OPT_FORCE_ALGO, P.FORCE_CELL_BINNED)
        sim.set_positions(pos)
        args.steps *= 10
    sim.advance(3)
    ms = sim.advance_timed(args.steps) / args.steps
    if args.clocks:                       # a long run with nvidia-smi sampling beside it: clock, power, throttle reasons
        import subprocess as sp, tempfile, time
        with tempfile.NamedTemporaryFile("w+", suffix=".csv") as f:
            mon = sp.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active", "--format=csv,noheader,nounits",
                            "-lms", "50"], stdout=f, stderr=sp.DEVNULL)
            time.sleep(0.3)
            ms_long = sim.advance_timed(args.clocks) / args.clocks
            mon.terminate(); mon.wait()
            f.seek(0)
            rows = [r.strip().split(", ") for r in f.read().strip().splitlines() if r.strip()]
        rows = rows[len(rows) // 3:]      # the loaded part
        mhz = sorted(float(r[0]) for r in rows)
        watt = sorted(float(r[1]) for r in rows)
        print("%s   long run: %.4f ms/step, SM clock median %.0f MHz (min %.0f), power median %.0f W, reasons %s" % (
            args.child, ms_long, mhz[len(mhz) // 2], mhz[0], watt[len(watt) // 2], sorted(set(r[2] for r in rows))), flush=True)
    c = sim.counters()
    ok = args.cells or int(c.filter_pairs) == int(sim.pairs_in_range())
    print("%s n=%d f%d: %.4f ms/step, filter audit %s" % (args.child, args.n, args.precision, ms, "ok" if ok else "FAILED"), flush=True)
    sim.close()
    return 0 if ok else 3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=262144)
    ap.add_argument("--precision", type=int, default=8)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--repeat", type=int, default=2)
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--cells", action="store_true", help="time the cell-binned path instead (no filter audit)")
    ap.add_argument("--clocks", type=int, default=0, help="steps of an extra run with nvidia-smi clock / power sampling")
    ap.add_argument("--child", default=None)
    ap.add_argument("libs", nargs="*")
    args = ap.parse_args()
    if args.child:
        return child(args)
    rc = 0
    for _ in range(args.repeat):
        for spec in args.libs:
            name, path = spec.split("=", 1)
            env = dict(os.environ, PSIM_LIBRARY=os.path.abspath(path))
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", name, "--n", str(args.n), "--precision", str(args.precision),
                                "--steps", str(args.steps), "--variant", str(args.variant), "--clocks", str(args.clocks)] + (["--cells"] if args.cells else []), env=env, timeout=600)
            rc |= r.returncode
    return rc


if __name__ == "__main__":
    sys.exit(main())
