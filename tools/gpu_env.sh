#!/bin/bash
# GPU visit for the env epilogue: all GPU tests, then short benches (plain step and env step).
# usage: gpu_env.sh tag
TAG=${1:-env}; O=gpurun_out/$TAG; mkdir -p $O
timeout 1200 python -m pytest tests -m gpu -q --timeout 900 -x > $O/pytest.log 2>&1; tail -5 $O/pytest.log
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d.get("env_step") or {}; print("%s %.2f M/s %.4f ms  med %.4f | av %.2f | env %.2f M/s %.4f ms | roll %.2f" % (sys.argv[1], d["value"]/1e6, d["ms_per_step"], d["ms_per_step_min_median_max"][1], d["with_angle_vel_epilogue"]["value"]/1e6, e.get("value",0)/1e6, e.get("ms_per_step",0), d["rollout"]["value"]/1e6) + (" | wrapped %.2f" % (e["with_episode_and_auto_reset_wrappers"]["value"]/1e6) if e else ""))'
for e in ant humanoid ant humanoid grasp fetch; do
  python bench.py --env $e --steps 100 --no-cpu --no-e2e 2> $O/bench_$e.err | tee $O/bench_$e.json | python -c "$P" $e
done
