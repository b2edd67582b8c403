"""Quick timing probe of one configuration: python tools/probe.py N PRECISION [allpairs|cells] [steps]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import particle_simulation_b200 as P
from particle_simulation_b200 import generate as G

n = int(sys.argv[1]); prec = int(sys.argv[2])
algo = sys.argv[3] if len(sys.argv) > 3 else "cells"
steps = int(sys.argv[4]) if len(sys.argv) > 4 and sys.argv[4].isdigit() else 100
ids, pos, vel, area = G.synthetic(n)
sim = P.Simulation(P.make_params(length_of_area=area), ids, pos, vel, prec)
sim.set_option(P.OPT_FORCE_ALGO, P.FORCE_CELL_BINNED if algo == "cells" else P.FORCE_ALL_PAIRS)
if algo.startswith("allpairs") and algo[-1].isdigit():
    sim.set_option(P.OPT_EXACT_DIFF_FILTER, int(algo[-1]))      # allpairs1/2/3: kernel variants
    sim.set_positions(pos.astype(P.native.rt(prec)))
sim.advance(30 if algo == "cells" else 3)
ms = sim.advance_timed(steps) / steps
c = sim.counters()
print(f"{algo} n={n} f{prec*8} ms/step={ms:.4f} steps/s={1e3/ms:.1f} pairs/s={n*(n-1)/ms*1e3:.3e} rebuilds={c.rebuilds}")
if "--brief" in sys.argv:
    sys.exit(0)
sim.set_option(P.OPT_KERNEL_TIMING, 1)
k0 = sim.counters(); ms2 = sim.advance_timed(steps) / steps; k1 = sim.counters()
fk = (k1.force_ms - k0.force_ms) / max(1, k1.force_timed - k0.force_timed)
print(f"   with per-launch events: ms/step={ms2:.4f} force kernel={fk:.4f} ms other={ms2-fk:.4f} ms launches/step={(k1.kernel_launches-k0.kernel_launches-1)/steps:.1f}")
