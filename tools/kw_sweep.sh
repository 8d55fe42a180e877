#!/bin/bash
# BASELINE.json configs[3] and configs[4] on one GPU: the HERRO-shaped workload and the k/w sweep on the chr20-shaped reads.
# One JSON line per run into gpurun_out/kw_sweep.jsonl (resident throughput + e2e; the CPU baseline leg is skipped).
out=gpurun_out/kw_sweep.jsonl
: > $out
python bench.py --workload herro --no-cpu-baseline --e2e-steps 1 --steps 3 --warmup 3 >> $out 2>> gpurun_out/kw_sweep.err
for k in 501 1001 2001 3001 5001; do
  for w in 50 100 200; do
    timeout 300 python bench.py --k $k --w $w --no-cpu-baseline --e2e-steps 1 --steps 3 --warmup 3 >> $out 2>> gpurun_out/kw_sweep.err || echo "{\"k\": $k, \"w\": $w, \"failed\": true}" >> $out
  done
done
