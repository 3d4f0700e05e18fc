#!/bin/bash
# Development aid: device-timed throughput against table height (per-tile fixed costs show at few rows).
for r in 200 500 1007 2000 5055 10122 35804; do
  echo "== rows $r"
  python tools/quick_bench.py 5.08e7 $r 3 2>&1 | tail -1
done
