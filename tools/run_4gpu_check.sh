python -m pytest tests/test_multi_rank.py -x -q -m gpu > gpurun_out/t4_multi.txt 2>&1
for n in 2 4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 3 --warmup 2 --no-cpu > gpurun_out/bench_${n}gpu.json 2> gpurun_out/bench_${n}gpu.err
done
