python -m pytest tests -x -q -m gpu -k "rule" > gpurun_out/t_rules.txt 2>&1
python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --trace > gpurun_out/bench_gather_new.json 2>gpurun_out/bench_gather_new.err
KMLS_RULES_GATHER=old python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --trace > gpurun_out/bench_gather_old.json 2>gpurun_out/bench_gather_old.err
