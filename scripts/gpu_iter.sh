#!/bin/bash
# kernel iteration: parity tests, a short bench line, one ncu --set full capture of the frame kernel
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 100 --warmup 5 --no-cpu --no-e2e > gpurun_out/bench_iter.json 2> gpurun_out/bench_iter.err || tail -5 gpurun_out/bench_iter.err
python - <<'PY'
import json
d=json.load(open("gpurun_out/bench_iter.json"))
r=d["roofline"]
print("ms/step", round(d["ms_per_step"],4), "G ent/s", round(d["value"]/1e9,2), "frame ms", round(r["kernel_ms"],4), "GB/s", round(r["achieved"]), "frac", round(r["frac"],3), r["kernel_ms_by_group"])
PY
if [ "${1:-}" = "ncu" ]; then bash scripts/gpu_prof.sh; fi
