#!/bin/bash
# threads-per-tile / batch-size sweep (compact lines) -> gpurun_out/<tag>/sweep.txt
TAG=${1:-sweep}; O=gpurun_out/$TAG; mkdir -p $O
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%.4g M/s  b2b %.4g M/s  %.4f ms  e2e %.3g M/s clk %s" % (d["value"]/1e6, d["value_back_to_back"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["clocks"]["sm_mhz"]))'
{
for e in ant humanoid; do
for t in 128 256 320 384 512; do echo -n "$e B=8192 T=$t: "; python bench.py --env $e --steps 30 --warmup 5 --threads $t --no-cpu | python -c "$P"; done
done
for b in 1024 32768 131072 524288; do for t in 128 256; do echo -n "ant B=$b T=$t: "; python bench.py --steps 20 --warmup 5 --threads $t --envs $b --no-cpu | python -c "$P"; done; done
for e in grasp fetch; do for t in 128 256 512; do echo -n "$e B=8192 T=$t: "; python bench.py --env $e --steps 20 --warmup 5 --threads $t --no-cpu | python -c "$P"; done; done
} > $O/sweep.txt 2>&1
cat $O/sweep.txt
