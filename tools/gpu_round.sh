#!/bin/bash
# One GPU-box visit: parity tests, parity report, smoke, bench, sweep, ncu.
# usage (from repo root): gpurun --timeout 1500 -- 'bash tools/gpu_round.sh [tag]'
TAG=${1:-r}
O=gpurun_out/$TAG
mkdir -p $O
nvidia-smi -L > $O/gpu.txt 2>&1; nproc >> $O/gpu.txt
python -c "import jax; print('jax', jax.__version__)" >> $O/gpu.txt 2>&1
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
timeout 300 python tools/parity_report.py > $O/parity.txt 2>&1
timeout 120 python __graft_entry__.py smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 300 python bench.py > $O/bench_ant.json 2> $O/bench_ant.err
for e in humanoid grasp fetch; do
  timeout 200 python bench.py --env $e --no-cpu --steps 50 > $O/bench_$e.json 2> $O/bench_$e.err
done
timeout 200 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.json 2>&1
CMD="python bench.py --steps 3 --warmup 3 --no-cpu"
$CMD > $O/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches.csv $CMD > $O/ncu1.log 2>&1
$CMD > $O/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:bx_step -s 25 -c 1 -o $O/prof $CMD > $O/ncu2.log 2>&1
tail -3 $O/pytest.log; cat $O/smoke.log | tail -2; cat $O/bench_ant.json
