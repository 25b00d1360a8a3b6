#!/bin/bash
# threads-per-tile sweep with the chunk-timed bench. usage: wsweep.sh tag "envs" "threads"
TAG=${1:-w}; O=gpurun_out/$TAG; mkdir -p $O
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%s T=%s %.2f M/s %.4f ms | env %.2f" % (sys.argv[1], sys.argv[2], d["value"]/1e6, d["ms_per_step"], (d.get("env_step") or {}).get("value",0)/1e6))'
for e in $2; do for t in $3; do
  python bench.py --env $e --steps 100 --no-cpu --no-e2e --threads $t 2>> $O/err.txt | python -c "$P" $e $t
done; done | tee $O/wsweep.txt
