python -m pytest tests/test_native_loop.py -x -q > gpurun_out/t_native.txt 2>&1
for sq in 100 50 25; do
  KMLS_SQUEEZE_PCT=$sq python bench.py --steps 3 --warmup 2 --no-cpu --trace > gpurun_out/bench_sq$sq.json 2>gpurun_out/bench_sq$sq.err
done
