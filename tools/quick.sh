#!/bin/bash
# quick GPU check: parity tests (all, no -x) + short benches. usage: quick.sh tag [pytest-args]
TAG=${1:-q}; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > $O/pytest.log 2>&1; tail -4 $O/pytest.log
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%.4g M/s  flushed %.4g M/s  %.4f ms  e2e %.3g M/s clk %s" % (d["value"]/1e6, d["value_l2_flushed"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["clocks"]["sm_mhz"]))'
{
for e in ant humanoid grasp fetch; do for t in 0 256; do echo -n "$e B=8192 T=$t: "; python bench.py --env $e --steps 30 --warmup 5 --threads $t --no-cpu | python -c "$P"; done; done
for t in 128 256; do echo -n "ant B=262144 T=$t: "; python bench.py --steps 10 --warmup 3 --threads $t --envs 262144 --no-cpu | python -c "$P"; done
} > $O/bench.txt 2>&1
cat $O/bench.txt
