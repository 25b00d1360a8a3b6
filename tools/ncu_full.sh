#!/bin/bash
# one ncu --set full capture of the step kernel (default config) + exports: ncu_full.sh tag [env]
TAG=${1:-n}; ENV=${2:-ant}; O=gpurun_out/$TAG; mkdir -p $O
CMD="python bench.py --env $ENV --steps 3 --warmup 3 --no-cpu --no-e2e"
$CMD > $O/plain.log 2>&1 || { echo "plain run failed"; tail -5 $O/plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:bx_step -s 25 -c 1 -o $O/prof $CMD > $O/ncu2.log 2>&1
ncu -i $O/prof.ncu-rep --page raw --csv > $O/raw.csv 2>/dev/null
ncu -i $O/prof.ncu-rep --page source --csv > $O/src.csv 2>/dev/null
ls -la $O; tail -2 $O/ncu2.log
