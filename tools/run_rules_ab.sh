python -m pytest tests -x -q -m gpu -k "rule" > gpurun_out/t_rules.txt 2>&1
python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --trace > gpurun_out/bench_rules_tma.json 2>gpurun_out/bench_rules_tma.err
KMLS_RULES_WRITE=r python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --trace > gpurun_out/bench_rules_reg.json 2>gpurun_out/bench_rules_reg.err
