#!/bin/bash
# ncu --set full of every hot kernel on c2 / c3 / c4s; the reports are digested ON the box (json + top source lines)
# and deleted, because gpurun brings back at most 64 MiB.   usage: bash tools/r02_ncu_all.sh <tag> [workloads] [kernels]
TAG=${1:-r02_x}; O=gpurun_out/$TAG; mkdir -p $O; WLS=${2:-"c2 c3 c4s"}; KS=${3:-"search_fwd_tile search_bwd wpsum_bwd wpsum_fwd"}
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem,power.draw,clocks_event_reasons.active --format=csv > $O/clocks_before.csv
for WL in $WLS; do
  python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline --graph off --blocks none > $O/plain_$WL.log 2>&1 || { echo "plain $WL failed"; continue; }
  for K in $KS; do
    R=/tmp/full_${K}_$WL
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:$K --launch-skip 3 --launch-count 1 -o $R -f \
      python bench.py --workload $WL --steps 1 --warmup 3 --no-cpu-baseline --graph off --blocks none > $O/ncu_${K}_$WL.log 2>&1
    echo "ncu $K $WL rc=$?"
    [ -f $R.ncu-rep ] || continue
    python tools/ncu_summary.py $O/ncu_${K}_$WL.json $R.ncu-rep > /dev/null
    python tools/ncu_lines.py $R.ncu-rep 30 > $O/lines_${K}_$WL.txt 2>/dev/null
    ncu -i $R.ncu-rep --page details 2>/dev/null | grep -E "Stall|stall|Warp Cycles Per|Eligible|Issued Warp|No Eligible|Registers Per|Theoretical Occ|Achieved Occ|Bank conflict|L2 Hit|L1/TEX Hit|Mem Busy|Max Bandwidth|Mem Pipes" > $O/details_${K}_$WL.txt
    ncu -i $R.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]; v=rows[2]
for i,k in enumerate(h):
    if k.startswith('smsp__average_warps_issue_stalled') and k.endswith('_per_issue_active.ratio') or k.startswith('smsp__average_warp_latency_issue_stalled'): print(k, v[i])
" > $O/stalls_${K}_$WL.txt
    if [ "$WL" = "c2" ] && [ "$K" = "search_fwd_tile" ]; then python tools/ncu_tile_phases.py $R.ncu-rep 204800 > $O/tile_phases_c2.txt; fi
    if [ "$WL" = "c3" ] && [ "$K" = "search_fwd_tile" ]; then python tools/ncu_tile_phases.py $R.ncu-rep 12582912 > $O/tile_phases_c3.txt; fi
    if [ "$WL" = "c4s" ] && [ "$K" = "search_fwd_tile" ]; then python tools/ncu_tile_phases.py $R.ncu-rep 50803200 > $O/tile_phases_c4s.txt; fi
    rm -f $R.ncu-rep
  done
done
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks.mem,power.draw,clocks_event_reasons.active --format=csv > $O/clocks_after.csv
du -sh $O
