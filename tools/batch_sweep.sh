#!/bin/bash
# BASELINE config 5: ant batch sweep 1K..1M envs on one GPU (throughput vs batch vs roofline).
# usage: batch_sweep.sh tag [env]
TAG=${1:-sweep}; ENV=${2:-ant}; O=gpurun_out/$TAG; mkdir -p $O
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d.get("env_step") or {}; print("%s B=%-8s %8.2f M env-steps/s  %9.4f ms/step  fp32 %.3f  hbm %.3f | env_step %.2f M/s | rollout10 %.2f M/s" % (sys.argv[1], sys.argv[2], d["value"]/1e6, d["ms_per_step"], d["roofline_fp32"]["frac"], d["roofline"]["frac"], e.get("value",0)/1e6, d["rollout"]["value"]/1e6))'
for B in ${3:-1024 4096 8192 16384 65536 262144 1048576}; do
  python bench.py --env $ENV --envs $B --steps 30 --warmup 5 --no-cpu --no-e2e 2> $O/err_$B.txt | tee $O/bench_$B.json | python -c "$P" $ENV $B
done | tee $O/sweep_$ENV.txt
