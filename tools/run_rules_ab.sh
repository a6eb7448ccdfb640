python -m pytest tests/test_native_loop.py -x -q > gpurun_out/t_native.txt 2>&1
python bench.py --steps 5 --warmup 3 --no-cpu --no-dense --trace > gpurun_out/bench_scan_new.json 2>gpurun_out/bench_scan_new.err
KMLS_SCAN=cub python bench.py --steps 5 --warmup 3 --no-cpu --no-dense --trace > gpurun_out/bench_scan_cub.json 2>gpurun_out/bench_scan_cub.err
