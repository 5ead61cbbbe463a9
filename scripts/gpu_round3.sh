#!/bin/bash
# round 3, one gpurun call on one GPU: the bench line, the ncu launch list, ncu --set full of the headline
# frame_kernel instantiation (16 Mi entities, 11 views) and of expand_lists_kernel (config 2), device times of configs 1, 2, 4
set -u
mkdir -p gpurun_out
python bench.py --steps 100 --warmup 5 > gpurun_out/r03_bench.json 2> gpurun_out/r03_bench.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r03_bench.err
python scripts/bench_configs.py --steps 50 > gpurun_out/r03_bench_configs.jsonl 2> gpurun_out/r03_bench_configs.err; echo "configs rc=$?"
PROF="python bench.py --steps 3 --warmup 3 --no-cpu --no-e2e"
$PROF > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r03_launches.csv $PROF > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:frame_kernel -s 3 -c 1 -o gpurun_out/r03_prof_frame -f $PROF > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
CFG2="python scripts/bench_configs.py --steps 3 --only 2"
$CFG2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:expand_lists_kernel -s 2 -c 1 -o gpurun_out/r03_prof_expand_lists -f $CFG2 > gpurun_out/ncu_full2.log 2>&1
echo "ncu expand_lists rc=$?"
ls -la gpurun_out | tail -12
