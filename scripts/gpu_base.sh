#!/bin/bash
# baseline check: GPU parity tests, the three side configs, the headline bench (no CPU leg)
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 300 python scripts/bench_configs.py --steps 50 > gpurun_out/bench_configs.jsonl 2> gpurun_out/bench_configs.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open("gpurun_out/bench_configs.jsonl"):
    d=json.loads(l); print(d["config"][:8], round(d["ms_per_step"],4), {k:round(v,4) for k,v in d["kernel_ms_by_group"].items()})
PY
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench rc=$?"
cat gpurun_out/bench_iter.json
