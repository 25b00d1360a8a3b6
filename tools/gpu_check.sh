#!/bin/bash
# quick regression visit: all GPU tests, ant + humanoid benches (plain / env / wrapped / rollout), smoke
python -m pytest tests -m gpu -q -x 2>&1 | tail -3
P='import json,sys; d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); e=d["env_step"]; print(sys.argv[1], "%.1f M | env %.1f | wrapped %.1f | roll %.1f" % (d["value"]/1e6, e["value"]/1e6, e["with_episode_and_auto_reset_wrappers"]["value"]/1e6, d["rollout"]["value"]/1e6))'
for e in ant humanoid; do
  python bench.py --env $e --no-cpu --no-e2e --steps 100 > gpurun_out/chk_$e.json 2> gpurun_out/chk_$e.err
  python -c "$P" gpurun_out/chk_$e.json
done
python __graft_entry__.py smoke | tail -1
