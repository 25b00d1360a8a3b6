#!/bin/bash
# A/B of programmatic dependent launch (BX_PDL=0/1), interleaved. usage: pdl_ab.sh tag
TAG=${1:-pdl}; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > $O/pytest.log 2>&1; tail -3 $O/pytest.log
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%s pdl=%s per-step-events %.2f M/s (%.4f ms) | one event pair %.2f M/s (%.4f ms) | b2b %.2f | env %.2f | rollout %.2f" % (sys.argv[1], sys.argv[2], d["value"]/1e6, d["ms_per_step"], d["value_one_event_pair"]/1e6, d["ms_per_step_one_event_pair"], d["value_back_to_back"]/1e6, (d.get("env_step") or {}).get("value",0)/1e6, d["rollout"]["value"]/1e6))'
for r in 1 2; do for e in ant humanoid; do for pdl in 0 1; do
  BX_PDL=$pdl python bench.py --env $e --steps 100 --no-cpu --no-e2e 2>> $O/err.txt | python -c "$P" $e $pdl
done; done; done | tee $O/ab.txt
