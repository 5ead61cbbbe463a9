#!/bin/bash
# ncu --set full of the wavefront frame kernel + expand on configs 2 and 4, and the tile timelines
set -u
mkdir -p gpurun_out
for c in 2 4; do
timeout 300 python scripts/trace_tiles.py $c > gpurun_out/trace_cfg$c.json 2> gpurun_out/trace_cfg$c.err; echo "trace cfg$c rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"frame_kernel|expand" -s 40 -c 2 -o gpurun_out/prof_cfg$c -f \
     python scripts/bench_configs.py --steps 30 --only $c > gpurun_out/ncu_cfg$c.log 2>&1; echo "ncu cfg$c rc=$?"
done
