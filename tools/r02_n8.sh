#!/bin/bash
# N ranks: the C2 line (no extra blocks) with the channels-last halo and with the planar one
TAG=${1:-r02_x}; N=${2:-8}; O=gpurun_out/$TAG; mkdir -p $O
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --blocks none; }
run > $O/bench_n${N}_cl.json 2> $O/bench_n${N}_cl.err; echo "bench N=$N (cl halo) rc=$?"; python benchsum.py < $O/bench_n${N}_cl.json
STNLS_HALO_CL=0 run > $O/bench_n${N}_planar.json 2> $O/bench_n${N}_planar.err; echo "bench N=$N (planar halo) rc=$?"; python benchsum.py < $O/bench_n${N}_planar.json
