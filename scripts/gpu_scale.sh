#!/bin/bash
# weak-scaling bench on the GPUs of one box: N = 8, 4 (the box has 8), each as torchrun would be launched by the driver
set -u
mkdir -p gpurun_out
for N in "$@"; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) \
     bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
  echo "N=$N rc=$?"
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_n$N.json").read().strip().splitlines()[-1])
    print("N=$N", round(d["value"]/1e9,2), "G ent/s", round(d["ms_per_step"],4), "ms/step; e2e", round(d["e2e"]["value"]/1e9,3), d["config"]["gather"], d["roofline"]["kernel_ms_by_group"])
except Exception as e:
    print("parse failed", e); print(open("gpurun_out/bench_n$N.err").read()[-1500:])
PY
done
