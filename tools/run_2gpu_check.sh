python -m pytest tests/test_native_loop.py -x -q > gpurun_out/t2_native.txt 2>&1
python -m pytest tests/test_multi_rank.py -x -q -m gpu > gpurun_out/t2_multi.txt 2>&1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$T --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 2 --no-cpu > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err
$T --master-port 29512 bench.py --gpus 2 --steps 3 --warmup 2 --no-cpu --partition playlists > gpurun_out/bench_2gpu_playlists.json 2> gpurun_out/bench_2gpu_playlists.err
python bench.py --steps 3 --warmup 2 --no-cpu --trace > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err
