#!/bin/bash
# The 8-GPU session: config 3 (strong scaling, candidate space split). "with5": also config 5 (auto: playlists sharded).
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
$T --master-port 29581 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu > gpurun_out/bench_8gpu.json 2> gpurun_out/bench_8gpu.err
if [ "$1" == "with5" ]; then
$T --master-port 29582 bench.py --gpus 8 --workload zipf_10Mx2M_s0.001 --steps 5 --warmup 2 --no-cpu > gpurun_out/r2_bench_z10m_8gpu.json 2> gpurun_out/r2_bench_z10m_8gpu.err
fi
nproc > gpurun_out/nproc8.txt
