#!/bin/bash
# BASELINE.json config 4: min_support sweep on the 1M x 100k Zipf set (one GPU).
# Usage (on the GPU box): bash tools/sweep.sh > gpurun_out/sweep.jsonl
# Rules are generated down to 0.001; at 0.0005 the rule table would exceed the 4e9-rule guard
# (the run would stop with an error, by design), so that point is mined without rules.
for s in 0.05 0.02 0.01 0.005 0.002 0.001; do
  python bench.py --workload zipf_1Mx100k_s0.001 --min-support $s --steps 2 --warmup 3 --no-cpu 2>> gpurun_out/sweep.log
done
python bench.py --workload zipf_1Mx100k_s0.001 --min-support 0.0005 --no-rules --steps 2 --warmup 3 --no-cpu 2>> gpurun_out/sweep.log
