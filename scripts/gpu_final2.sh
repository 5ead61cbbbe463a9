#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 25 python -m pytest -x -q -m gpu tests/test_shim_gpu.py > gpurun_out/pytest_shim.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/pytest_shim.log | cut -c1-300
timeout 25 python bench.py --shim-e2e 3 --entities 4194304 --e2e-steps 5 > gpurun_out/shim_e2e_4m.json 2> gpurun_out/shim_e2e_4m.err; echo "shim rc=$?"
cat gpurun_out/shim_e2e_4m.json | cut -c1-400; tail -2 gpurun_out/shim_e2e_4m.err | cut -c1-300
