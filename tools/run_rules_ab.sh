python -m pytest tests -x -q -m gpu -k "rule or artifact" > gpurun_out/t_rules.txt 2>&1
python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --trace > gpurun_out/bench_subs_new.json 2>gpurun_out/bench_subs_new.err
KMLS_RULES_SUBSETS=old python bench.py --steps 3 --warmup 2 --no-cpu --no-dense > gpurun_out/bench_subs_old.json 2>gpurun_out/bench_subs_old.err
