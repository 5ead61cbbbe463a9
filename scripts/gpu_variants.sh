#!/bin/bash
# bench every prebuilt library variant under build/variants/ (kernel tuning sweeps)
cp pmtech_b200/lib/libpmtech_b200.so /tmp/keep.so
for f in build/variants/*.so; do
  cp $f pmtech_b200/lib/libpmtech_b200.so
  echo -n "$(basename $f) "
  python bench.py --steps 100 --warmup 5 --no-cpu --no-e2e | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), round(d['roofline']['kernel_ms'],4), round(d['roofline']['frac'],4))"
done
cp /tmp/keep.so pmtech_b200/lib/libpmtech_b200.so
