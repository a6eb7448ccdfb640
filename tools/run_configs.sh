#!/bin/bash
# Builder-run evidence for BASELINE.json's configs 1, 2, 4 and 5 on ONE B200 (config 3 is bench.py's default).
# Every line carries the parity object (oracle result committed in tests/golden/).
out=gpurun_out
for w in mpd_standin_seed1_s0.05 mpd_standin_seed2_s0.05; do
  python bench.py --workload $w --steps 20 --warmup 5 --no-dense > $out/r2_bench_$w.json 2> $out/r2_bench_$w.err
done
: > $out/r2_sweep_z1m.jsonl
for s in 0.05 0.02 0.01 0.005 0.002 0.001; do
  python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --min-support $s >> $out/r2_sweep_z1m.jsonl 2>> $out/r2_sweep.err
done
python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --min-support 0.0005 --no-rules >> $out/r2_sweep_z1m.jsonl 2>> $out/r2_sweep.err
python bench.py --workload dense_200kx4k_s0.02 --steps 5 --warmup 3 --no-cpu --no-dense --trace > $out/r2_bench_dense.json 2> $out/r2_bench_dense.err
if [ "$1" == "with5" ]; then
python bench.py --workload zipf_10Mx2M_s0.001 --steps 3 --warmup 2 --no-cpu --no-dense > $out/r2_bench_z10m_1gpu.json 2> $out/r2_bench_z10m_1gpu.err
fi
