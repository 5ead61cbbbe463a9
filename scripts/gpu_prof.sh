#!/bin/bash
# one ncu --set full capture of the frame kernel (after a plain run of the same command exited 0)
set -u
mkdir -p gpurun_out
PROF="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e"
$PROF > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:frame_kernel -s 3 -c 1 -o gpurun_out/prof_frame -f $PROF > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
