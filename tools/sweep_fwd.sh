#!/bin/bash
# sweep the forward-table knobs (groups per batch budget, warps per CTA) on the bench workloads
O=gpurun_out/${1:-sweep}; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for wl in c2 c3 c4s; do
 for bud in 6144 9216 14336; do
  for w in 4 8; do
   echo "== $wl budget=$bud warps=$w"
   STNLS_FWD_BUDGET=$bud STNLS_FWD_WARPS=$w timeout 300 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline 2>&1 | python benchsum.py
  done
 done
done 2>&1 | tee $O/sweep.txt
