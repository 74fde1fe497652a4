#!/bin/bash
# ncu --set full of the search forward kernel on a workload: tools/ncu_fwd.sh <tag> <workload> [kernel regex]
O=gpurun_out/$1; mkdir -p $O; WL=${2:-c2}; K=${3:-search_fwd}
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K --launch-skip 3 --launch-count 1 \
   -o $O/full_${K}_$WL -f python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline > $O/ncu_full_${K}_$WL.log 2>&1; echo "ncu rc=$?"
