#!/bin/bash
# round 2: parity tests, the other-config timings, a short bench line
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 300 python scripts/bench_configs.py --steps 50 > gpurun_out/bench_configs.jsonl 2> gpurun_out/bench_configs.err; echo "configs rc=$?"
cat gpurun_out/bench_configs.jsonl; tail -3 gpurun_out/bench_configs.err
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench rc=$?"
cat gpurun_out/bench_iter.json; tail -3 gpurun_out/bench_iter.err
