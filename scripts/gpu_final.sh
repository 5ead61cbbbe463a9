#!/bin/bash
# last check of the round on one GPU: the tests whose code changed since the last full green run, then a short bench
set -u
mkdir -p gpurun_out
timeout 55 python -m pytest -x -q -m gpu tests/test_shim_gpu.py tests/test_c_consumer.py tests/test_parity_gpu.py -k "shim or group or consumer or instance or row_exchange or peer_push" > gpurun_out/pytest_final.log 2>&1; echo "pytest rc=$?"
tail -6 gpurun_out/pytest_final.log | cut -c1-400
timeout 75 python bench.py --steps 20 --warmup 3 --no-cpu --entities 4194304 > gpurun_out/bench_final.json 2> gpurun_out/bench_final.err; echo "bench rc=$?"
tail -c 6000 gpurun_out/bench_final.json
tail -3 gpurun_out/bench_final.err | cut -c1-300
