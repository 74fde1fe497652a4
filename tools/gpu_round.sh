#!/bin/bash
# One GPU-box pass: parity tests, bench lines, ncu launch list, ncu --set full of the top kernel.
# usage (via gpurun): bash tools/gpu_round.sh <tag>
TAG=${1:-r01_x}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > $O/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_c2.json 2> $O/bench_c2.err; echo "bench c2 rc=$?"
timeout 600 python bench.py --workload c3 --steps 5 --warmup 3 --no-cpu-baseline > $O/bench_c3.json 2> $O/bench_c3.err; echo "bench c3 rc=$?"
timeout 600 python bench.py --workload c4s --steps 3 --warmup 3 --no-cpu-baseline > $O/bench_c4s.json 2> $O/bench_c4s.err; echo "bench c4s rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref_c2.json 2> $O/bench_ref_c2.err; echo "bench ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c2.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu-baseline --graph off > $O/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
for K in search_fwd wpsum_bwd search_bwd wpsum_fwd; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$K --launch-skip 3 --launch-count 1 \
   -o $O/full_${K}_c2 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --graph off > $O/ncu_full_$K.log 2>&1; echo "ncu full $K rc=$?"
done
python benchsum.py < $O/bench_c2.json; python benchsum.py < $O/bench_c3.json; python benchsum.py < $O/bench_c4s.json
