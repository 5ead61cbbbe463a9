#!/bin/bash
for cfg in "2 0 2" "2 1 2" "3 1 3"; do
  echo "== stream waits: $cfg"; timeout 30 python scripts/debug_group.py $cfg 2>&1 | grep -v "^rank" | tail -4
done
echo "== poll kernels: 3 1 3"; timeout 40 python scripts/debug_group.py 3 1 3 2>&1 | grep -v "^rank" | tail -4
