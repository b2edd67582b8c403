import sys, time; sys.path.insert(0,'.')
import numpy as np
import particle_simulation_b200 as P
from particle_simulation_b200 import generate as G
for prec in (8,4):
    print('peak', prec, P.native.measure_fma_peak(prec))
for n in (65536,):
    ids,pos,vel,area=G.synthetic(n)
    p=P.make_params(length_of_area=area)
    for prec in (8,4):
        for exact in (0,1):
            sim=P.Simulation(p,ids,pos,vel,prec)
            sim.set_option(P.OPT_EXACT_DIFF_FILTER, exact)
            sim.advance(3)
            ms=sim.advance_timed(20)/20
            pin=sim.pairs_in_range()
            print(f'n={n} prec={prec} exact={exact} ms/step={ms:.4f} pairs/s={n*(n-1)/ms*1e3:.3e} P_in={pin} rebuilds={sim.counters().rebuilds}')
            sim.close()
