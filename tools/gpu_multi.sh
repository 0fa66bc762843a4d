#!/bin/bash
# usage: gpurun --gpus N -- bash tools/gpu_multi.sh N   (multi-GPU parity check + bench under torchrun)
mkdir -p gpurun_out
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py > gpurun_out/multi_check_n$N.json 2> gpurun_out/multi_check_n$N.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
tail -n 3 gpurun_out/multi_check_n$N.err; tail -n 3 gpurun_out/bench_n$N.err
