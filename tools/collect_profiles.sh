#!/bin/bash
# copy the evidence of a tools/gpu_round.sh pass from gpurun_out/<tag>/ into profiles/ and refresh traffic.json
TAG=$1; S=gpurun_out/$TAG
cp $S/launches_c2.csv profiles/${TAG}_launches_c2.csv
for w in c2 c3 c4s; do tail -1 $S/bench_$w.json > profiles/${TAG}_bench_$w.json; done
tail -1 $S/bench_ref_c2.json > profiles/${TAG}_bench_ref_c2.json
tail -4 $S/pytest_gpu.log > profiles/${TAG}_pytest_gpu.txt
python tools/ncu_summary.py profiles/${TAG}_ncu_full_c2.json $S/full_*.ncu-rep
python tools/ncu_phases.py $S/full_search_fwd_c2.ncu-rep search_fwd_tab_kernel.cuh:54-82=lattice_value search_fwd_tab_kernel.cuh:84-164=setup search_fwd_tab_kernel.cuh:165-204=anchor_pass search_fwd_tab_kernel.cuh:205-268=axis_tables search_fwd_tab_kernel.cuh:269-302=lattice search_fwd_tab_kernel.cuh:303-575=candidates search_fwd_tab_kernel.cuh:576-607=select search_fwd_tab_kernel.cuh:608-660=write > profiles/${TAG}_search_fwd_phases.txt
python tools/ncu_smem.py $S/full_search_fwd_c2.ncu-rep 12 > profiles/${TAG}_search_fwd_smem.txt
python - <<PY
import json
s=json.load(open('profiles/${TAG}_ncu_full_c2.json'))
m={'full_search_fwd_c2.ncu-rep':'search_forward','full_search_bwd_c2.ncu-rep':'search_backward','full_wpsum_fwd_c2.ncu-rep':'wpsum_forward','full_wpsum_bwd_c2.ncu-rep':'wpsum_backward'}
out={"_source":"ncu --set full --clock-control none, one launch per kernel, dram__bytes_read.sum + dram__bytes_write.sum (profiles/${TAG}_ncu_full_c2.json)","c2":{}}
for k,v in s.items(): out["c2"][m[k]]=v[0]["traffic_bytes"]
json.dump(out,open('profiles/traffic.json','w'),indent=1); print(out)
PY
