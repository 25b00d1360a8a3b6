#!/bin/bash
# e2e (host path) vs the stagger rate: e2e_sweep.sh tag env
O=gpurun_out/${1:-e2e}; mkdir -p $O; ENV=${2:-ant}
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("e2e %.2f M/s" % (d["e2e"]["value"]/1e6))'
for g in ${GBPS:-0 25 35 45 60}; do
  echo -n "$ENV stagger_gbps=$g: "; BX_HOST_STAGGER_GBPS=$g python bench.py --env $ENV --steps 30 --no-cpu | python -c "$P"
done | tee $O/e2e_$ENV.txt
