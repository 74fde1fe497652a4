#!/bin/bash
# bench workloads only (no tests); env knobs pass through
for wl in ${@:-c2 c3 c4s}; do echo "== $wl"; timeout 300 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python benchsum.py; done
