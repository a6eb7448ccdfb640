python -m pytest tests -x -q -m gpu -k "rule" > gpurun_out/t_rules.txt 2>&1
python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --trace > gpurun_out/bench_rules_w2.json 2>gpurun_out/bench_rules_w2.err
B="python bench.py --steps 1 --warmup 0 --no-cpu --no-parity --no-dense --mine-ws-gb 28 --rules-ws-gb 84"
ncu --set full --clock-control none --import-source on -k regex:k_rules_write_mlp --launch-skip 2 -c 1 -f -o gpurun_out/r2_rules_write_mlp2 $B > gpurun_out/ncu_rules_write_mlp2.log 2>&1
