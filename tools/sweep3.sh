#!/bin/bash
TAG=${1:-sweep}; O=gpurun_out/$TAG; mkdir -p $O
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%.4g M/s  b2b %.4g M/s  %.4f ms" % (d["value"]/1e6, d["value_back_to_back"]/1e6, d["ms_per_step"]))'
{
for e in ant humanoid grasp fetch; do
for t in 256 320 384 512; do echo -n "$e B=8192 T=$t: "; python bench.py --env $e --steps 30 --warmup 5 --threads $t --no-cpu --no-e2e | python -c "$P"; done
done
for e in ant humanoid; do for t in 256 320; do echo -n "$e B=262144 T=$t: "; python bench.py --env $e --steps 10 --warmup 3 --threads $t --envs 262144 --no-cpu --no-e2e | python -c "$P"; done; done
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
