KMLS_AND_HALF=1 python -m pytest tests/test_native_loop.py -x -q > gpurun_out/t_native.txt 2>&1
python bench.py --steps 3 --warmup 2 --no-cpu --no-dense --trace > gpurun_out/bench_half_auto.json 2>gpurun_out/bench_half_auto.err
KMLS_AND_HALF=0 python bench.py --steps 3 --warmup 2 --no-cpu --no-dense > gpurun_out/bench_half_off.json 2>gpurun_out/bench_half_off.err
