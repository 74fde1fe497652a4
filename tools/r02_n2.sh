#!/bin/bash
# 2 GPUs: sharded == single GPU for every halo path, then the C2 line with the channels-last exchange and with the planar one
TAG=${1:-r02_w}; N=${2:-2}; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests/test_parallel_gpu.py -q -x > $O/pytest_parallel.txt 2>&1; echo "parallel tests rc=$?"; tail -12 $O/pytest_parallel.txt
run() { timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline --blocks none; }
run > $O/bench_n${N}_cl.json 2> $O/bench_n${N}_cl.err; echo "bench N=$N (cl halo) rc=$?"; python benchsum.py < $O/bench_n${N}_cl.json
STNLS_HALO_CL=0 run > $O/bench_n${N}_planar.json 2> $O/bench_n${N}_planar.err; echo "bench N=$N (planar halo) rc=$?"; python benchsum.py < $O/bench_n${N}_planar.json
