#!/bin/bash
# A/B of three tiles per SM (BX_THREE_TILES=0/1) over batch sizes. usage: three_ab.sh tag "envs" "batches"
TAG=${1:-three}; O=gpurun_out/$TAG; mkdir -p $O
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%s B=%s three=%s %.2f M/s %.4f ms | env %.2f | roll %.2f" % (sys.argv[1], sys.argv[2], sys.argv[3], d["value"]/1e6, d["ms_per_step"], (d.get("env_step") or {}).get("value",0)/1e6, d["rollout"]["value"]/1e6))'
for e in $2; do for B in $3; do for t in 0 1; do
  BX_THREE_TILES=$t python bench.py --env $e --envs $B --steps 30 --warmup 5 --no-cpu --no-e2e 2>> $O/err.txt | python -c "$P" $e $B $t
done; done; done | tee $O/ab.txt
