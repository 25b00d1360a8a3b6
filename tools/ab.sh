#!/bin/bash
# A/B of two library builds, interleaved N times: ab.sh LIB_A LIB_B [env] [n] [extra bench args]
A=$1; B=$2; ENV=${3:-ant}; N=${4:-4}; shift 4
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%s %.2f M/s %.4f ms  med %.4f" % (sys.argv[1], d["value"]/1e6, d["ms_per_step"], d["ms_per_step_min_median_max"][1]))'
for i in $(seq $N); do
  BRAX_B200_LIB=$A python bench.py --env $ENV --steps 100 --no-cpu --no-e2e "$@" | python -c "$P" A
  BRAX_B200_LIB=$B python bench.py --env $ENV --steps 100 --no-cpu --no-e2e "$@" | python -c "$P" B
done
