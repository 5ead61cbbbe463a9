#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 200 python scripts/trace_tiles.py 2 > gpurun_out/trace_cfg2.json 2> gpurun_out/trace_cfg2.err; echo "trace rc=$?"; tail -2 gpurun_out/trace_cfg2.err
timeout 200 python scripts/trace_tiles.py 4 > gpurun_out/trace_cfg4.json 2> gpurun_out/trace_cfg4.err; echo "trace rc=$?"
timeout 300 python scripts/bench_configs.py --steps 50 > gpurun_out/bench_configs.jsonl 2> gpurun_out/bench_configs.err; echo "configs rc=$?"
cat gpurun_out/bench_configs.jsonl; tail -3 gpurun_out/bench_configs.err
timeout 300 python scripts/bench_configs.py --steps 50 --no-tail > gpurun_out/bench_configs_notail.jsonl 2> gpurun_out/bench_configs_notail.err; echo "configs no-tail rc=$?"
cat gpurun_out/bench_configs_notail.jsonl
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_iter.json"))
r=d["roofline"]
print("ms/step", round(d["ms_per_step"],4), "G ent/s", round(d["value"]/1e9,2), "frame ms", round(r["kernel_ms"],4), "frac", round(r["frac"],3), r["kernel_ms_by_group"])
for c in (2,4):
    t=json.load(open(f"gpurun_out/trace_cfg{c}.json"))
    print(c, {k:v for k,v in t.items() if k!="first_64_tiles"})
PY
