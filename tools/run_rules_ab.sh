python -m pytest tests -x -q -m gpu -k "rule or artifact or multi" > gpurun_out/t_rules.txt 2>&1
python bench.py --steps 5 --warmup 3 --no-cpu --no-dense --trace > gpurun_out/bench_trie.json 2>gpurun_out/bench_trie.err
