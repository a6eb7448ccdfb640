#!/bin/bash
# ncu evidence for the default path (1 GPU): a launch list (shares) and one --set full capture
# per hot kernel. Run on the GPU box AFTER the same bench command has exited 0 without ncu.
# Workspaces are pre-sized so that no call is repeated and launch indices are deterministic.
set -x
B="python bench.py --steps 1 --warmup 0 --no-cpu --no-parity --mine-ws-gb 28 --rules-ws-gb 84"
$B > gpurun_out/ncu_plain.json 2> gpurun_out/ncu_plain.err || exit 1
if [ "$1" != "quick" ]; then
ncu --metrics gpu__time_duration.sum --clock-control none -c 320 --csv --log-file gpurun_out/r2_launches.csv $B > gpurun_out/ncu_l.log 2>&1
fi
if [ "$1" = "pairs" ]; then   # the two level-2 kernels of the hybrid only
cap2() { ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $3 -c $4 -f -o gpurun_out/r2_$1 $B > gpurun_out/ncu_$1.log 2>&1; }
cap2 pair_csr_cold k_pair_counts_csr 0 1
cap2 pair_gemm_block k_pair_gemm_i8_ts 0 1
exit 0
fi
cap() {  # name, kernel regex, launches to skip, launches to capture
  ncu --set full --clock-control none --import-source on -k regex:$2 --launch-skip $3 -c $4 -f -o gpurun_out/r2_$1 $B > gpurun_out/ncu_$1.log 2>&1
}
cap count_project k_count_project 0 2      # levels 3 and 4 of the timed step
cap count_and k_count_and 4 1              # level 7 (levels 3.. = 11 launches per step)
if [ "$1" != "quick" ]; then
cap pair_csr k_pair_counts_csr 0 1
cap rules_gather k_rules_gather 5 1        # level 7 (levels 2.. = 12 launches per step)
cap rules_write k_rules_write 6 1          # level 8
fi
