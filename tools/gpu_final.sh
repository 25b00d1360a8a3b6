#!/bin/bash
# Round-end GPU visit: everything gpu_round.sh does, plus ncu launch lists of the env / reset path.
# usage: gpurun --timeout 2400 -- 'bash tools/gpu_final.sh tag'
TAG=${1:-final}; O=gpurun_out/$TAG
bash tools/gpu_round.sh $TAG
for e in ant fetch; do
  python tools/mini_env_step.py $e > $O/mini_$e.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $O/launches_env_$e.csv \
    python tools/mini_env_step.py $e > $O/ncu_env_$e.log 2>&1
  tail -1 $O/mini_$e.log
done
ncu -i $O/prof.ncu-rep --page raw --csv > $O/raw.csv 2>/dev/null
