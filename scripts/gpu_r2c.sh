#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for v in "" "--no-tail"; do
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
for c in 2 4; do
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"frame_kernel|expand_kernel" -s 40 -c 2 -o gpurun_out/prof_cfg$c -f \
     python scripts/bench_configs.py --steps 30 --only $c > gpurun_out/ncu_cfg$c.log 2>&1; echo "ncu cfg$c rc=$?"
done
