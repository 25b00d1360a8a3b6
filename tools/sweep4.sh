#!/bin/bash
# contact-section x warps sweep, interleaved twice: sweep4.sh tag
O=gpurun_out/${1:-s4}; mkdir -p $O
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%.2f M/s %.4f ms" % (d["value"]/1e6, d["ms_per_step"]))'
for rep in 1 2; do
for e in ant humanoid; do for cs in 1 2; do for t in 320 384; do
  echo -n "$e section=$cs T=$t: "; python bench.py --env $e --steps 100 --threads $t --contact-section $cs --no-cpu --no-e2e | python -c "$P"
done; done; done
for e in grasp fetch ant_spring humanoid_spring; do for t in 256 320 384; do
  echo -n "$e T=$t: "; python bench.py --env $e --steps 100 --threads $t --no-cpu --no-e2e | python -c "$P"
done; done
done | tee $O/sweep4.txt
