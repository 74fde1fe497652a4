#!/bin/bash
# backward kernels A/B: parity subset first, then per-kernel times on c2 / c4s with and without an env knob
TAG=${1:-r02_t}; KNOB=${2:-STNLS_NO_FRAMEGRID}; WLS=${3:-"c2 c4s"}; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x -k "named or parity or guard or abi or bwd or backward or grad" > $O/pytest_bwd.log 2>&1; echo "pytest rc=$?"; tail -3 $O/pytest_bwd.log
for wl in $WLS; do
  timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --blocks none > $O/bench_$wl.json 2> $O/bench_$wl.err; echo "bench $wl rc=$?"; python benchsum.py < $O/bench_$wl.json
  env $KNOB=1 timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --blocks none > $O/bench_${wl}_knob.json 2> $O/bench_${wl}_knob.err; echo "bench $wl ($KNOB=1) rc=$?"; python benchsum.py < $O/bench_${wl}_knob.json
done
