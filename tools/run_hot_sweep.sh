# level-2 hybrid (hot rows' block on the tensor cores): parity tests, then the bench at several block sizes
python -m pytest tests/test_native_loop.py -x -q -k "hybrid or cold_plus_block" > gpurun_out/t_hybrid.txt 2>&1
for h in -1 0 512 1024 2048; do
  KMLS_PAIR_HOT_ROWS=$h python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --trace > gpurun_out/bench_hot$h.json 2>gpurun_out/bench_hot$h.err
done
