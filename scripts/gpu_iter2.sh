#!/bin/bash
# iteration check: GPU parity tests, the three side configs (default and forced one-launch tail), headline bench
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for v in "" "--force-tail" "--no-tail"; do
  timeout 300 python scripts/bench_configs.py --steps 50 $v > gpurun_out/bench_configs$v.jsonl 2> gpurun_out/bench_configs$v.err; echo "configs $v rc=$?"
  python - "$v" <<'PY'
import json,sys
for l in open(f"gpurun_out/bench_configs{sys.argv[1]}.jsonl"):
    d=json.loads(l); print(sys.argv[1], d["config"][:8], round(d["ms_per_step"],4), {k:round(v,4) for k,v in d["kernel_ms_by_group"].items()})
PY
done
timeout 300 python bench.py --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_iter.json"))
r=d["roofline"]
print("ms/step", round(d["ms_per_step"],4), "G ent/s", round(d["value"]/1e9,2), "frame ms", round(r["kernel_ms"],4), "frac", round(r["frac"],3), r["kernel_ms_by_group"])
PY
