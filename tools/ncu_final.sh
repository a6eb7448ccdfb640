#!/bin/bash
# Final evidence of the round for the kernels that changed last: launch list of the default path and one
# --set full capture of the four-slots-per-lane gather (level 7). The other captures (count_project,
# count_and, pair_csr_cold, pair_gemm_block, rules_write_mlp) were taken with the kernels as they are now.
B="python bench.py --steps 1 --warmup 0 --no-cpu --no-parity --no-dense --mine-ws-gb 28 --rules-ws-gb 84"
$B > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 320 --csv --log-file gpurun_out/r2_launches.csv $B > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_rules_gather_v4 --launch-skip 3 -c 1 -f -o gpurun_out/r2_rules_gather_v4 $B > gpurun_out/ncu_rules_gather_v4.log 2>&1
