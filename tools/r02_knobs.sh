#!/bin/bash
# A/B of env knobs on one workload each: "<workload>:<ENV=val>[,<ENV=val>]" ...; baseline first
TAG=${1:-r02_y}; shift; O=gpurun_out/$TAG; mkdir -p $O
for spec in "$@"; do
  wl=${spec%%:*}; envs=${spec#*:}; [ "$envs" = "$spec" ] && envs=""
  name=$(echo "${wl}_${envs}" | tr -c 'A-Za-z0-9_\n' '_')
  env $(echo $envs | tr ',' ' ') timeout 300 python bench.py --workload $wl --steps 10 --warmup 3 --no-cpu-baseline --blocks none > $O/$name.json 2> $O/$name.err
  echo "== $wl [$envs] rc=$?"; python benchsum.py < $O/$name.json
done
