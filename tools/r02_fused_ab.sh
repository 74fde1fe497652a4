#!/bin/bash
TAG=${1:-r02_u}; O=gpurun_out/$TAG; mkdir -p $O
timeout 600 python -m pytest tests/test_fused_step_gpu.py -q -x > $O/pytest_fused.log 2>&1; echo "fused tests rc=$?"; tail -15 $O/pytest_fused.log
timeout 900 python -m pytest tests -m gpu -q -x > $O/pytest_gpu.txt 2>&1; echo "suite rc=$?"; tail -3 $O/pytest_gpu.txt
for wl in c2 c4s; do
  timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --blocks none > $O/bench_$wl.json 2> $O/bench_$wl.err; echo "bench $wl rc=$?"; python benchsum.py < $O/bench_$wl.json
  STNLS_FUSED_STEP=0 timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --blocks none > $O/bench_${wl}_unfused.json 2> $O/bench_${wl}_unfused.err; echo "bench $wl (module chain) rc=$?"; python benchsum.py < $O/bench_${wl}_unfused.json
done
