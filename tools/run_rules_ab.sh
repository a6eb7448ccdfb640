python -m pytest tests -x -q -m gpu > gpurun_out/t_full.txt 2>&1
python bench.py --steps 5 --warmup 3 --no-cpu --no-dense --trace > gpurun_out/bench_tid.json 2>gpurun_out/bench_tid.err
