#!/bin/bash
# threads-per-tile sweep at the default contact section: sweep5.sh tag envs...
O=gpurun_out/${1:-s5}; mkdir -p $O; shift
P='import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print("%.2f M/s %.4f ms" % (d["value"]/1e6, d["ms_per_step"]))'
for e in "$@"; do for t in 256 320 384 512; do
  echo -n "$e T=$t: "; python bench.py --env $e --steps 100 --threads $t --no-cpu --no-e2e | python -c "$P"
done; done | tee $O/sweep5.txt
