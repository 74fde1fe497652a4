#!/bin/bash
# parity tests + the three bench workloads, short
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for wl in ${@:-c2 c3 c4s}; do echo "== $wl"; timeout 300 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline 2>&1 | python benchsum.py; done
