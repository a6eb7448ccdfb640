T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$T --nproc-per-node 4 --master-port 29541 bench.py --gpus 4 --steps 5 --warmup 3 --no-cpu > gpurun_out/bench_4gpu.json 2> gpurun_out/bench_4gpu.err
python -m pytest tests/test_multi_rank.py -x -q -m gpu > gpurun_out/t4_multi.txt 2>&1
