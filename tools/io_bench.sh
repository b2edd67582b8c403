#!/bin/bash
# Frame pipeline timing: simu_b200 on the N=262,144 lattice, a frame every 10 steps,
# writer thread against --sync-io.  Run from the repo root on a GPU box; scratch files go to gpurun_out/io_bench/.
set -e
mkdir -p gpurun_out/io_bench
python - <<'PY'
import numpy as np
from particle_simulation_b200 import generate as G
ids,pos,vel,area=G.synthetic(262144)
with open("gpurun_out/io_bench/big.dat","w") as f:
    for i,p,v in zip(ids,pos,vel): f.write("%d %.4f %.4f %.4f %.4f %.4f %.4f\n"%(i,*p,*v))
open("gpurun_out/io_bench/big.nml","w").write("&ENVIRONMENTPARAMS\n NB_OF_THREADS=1,\n SIGMA=1.0,\n EPSILON=1.0,\n MASS=1.0,\n R_CUT=2.5,\n LENGTH_OF_AREA=%g,%g,%g,\n LENGTH_OF_CELL=3.0,3.0,3.0,\n DELTA_T=0.001,\n WRITE_TIME_STEP=10,\n MAX_TIME_STEP=400\n /\n"%tuple(area))
PY
for mode in "" "--sync-io"; do
  for rep in 1 2; do
    particle_simulation_b200/_build/simu_b200 gpurun_out/io_bench/big.nml gpurun_out/io_bench/big.dat gpurun_out/io_bench/big.output --precision 4 $mode | grep -E "elapsed|frames" | tr '\n' ' '; echo " mode=[$mode]"
  done
done
ls -la gpurun_out/io_bench/big.output
rm -f gpurun_out/io_bench/big.output gpurun_out/io_bench/big.dat
