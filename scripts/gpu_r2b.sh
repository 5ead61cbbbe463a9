#!/bin/bash
# ncu --set full of the wavefront form (configs 2 and 4, separate launches)
set -u
mkdir -p gpurun_out
for c in 4 2; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:frame_kernel -s 20 -c 1 -o gpurun_out/prof_cfg$c -f \
     python scripts/bench_configs.py --steps 30 --only $c --no-tail > gpurun_out/ncu_cfg$c.log 2>&1
  echo "ncu cfg$c rc=$?"
done
