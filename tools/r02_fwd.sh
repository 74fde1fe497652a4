#!/bin/bash
# tile-kernel bring-up: parity on the named configs, then timing of the three workloads with and without it
TAG=${1:-r02_c}; O=gpurun_out/$TAG; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q -x -k "named or full_size or ref_parity" > $O/pytest_fwd.log 2>&1; echo "pytest rc=$?"; tail -5 $O/pytest_fwd.log
for wl in c2 c3 c4s; do
  timeout 300 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_$wl.json 2> $O/bench_$wl.err; echo "bench $wl rc=$?"; python benchsum.py < $O/bench_$wl.json
  STNLS_FWD_TILE=0 timeout 300 python bench.py --workload $wl --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_${wl}_notile.json 2> $O/bench_${wl}_notile.err; echo "bench $wl (table kernel) rc=$?"; python benchsum.py < $O/bench_${wl}_notile.json
done
