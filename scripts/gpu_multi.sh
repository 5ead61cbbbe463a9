#!/bin/bash
# multi-GPU: the exchange check under pytest, then the bench at this call's GPU count (and the modes)
set -u
N=${1:-2}
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_multi_gpu.py -m gpu -x -q > gpurun_out/pytest_multi_gpu_n$N.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_multi_gpu_n$N.log
for mode in ${MODES:-root counts all}; do
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 100 --warmup 5 --exchange $mode ${2:-} > gpurun_out/bench_n${N}_$mode.json 2> gpurun_out/bench_n${N}_$mode.err; echo "bench $mode rc=$?"
python - $N $mode <<'PY'
import json,sys
try:
    d=json.load(open(f"gpurun_out/bench_n{sys.argv[1]}_{sys.argv[2]}.json"))
    print(sys.argv[2], "ms/step", round(d["ms_per_step"],4), "G ent/s", round(d["value"]/1e9,2), "frame ms", round(d["roofline"]["kernel_ms"],4), d["config"].get("exchange"), "e2e", d["e2e"] and round(d["e2e"]["value"]/1e9,3))
except Exception as e:
    print("no line", e); print(open(f"gpurun_out/bench_n{sys.argv[1]}_{sys.argv[2]}.err").read()[-2000:])
PY
done
