#!/bin/bash
# usage: gpurun --gpus N --timeout 900 -- bash tools/gpu_multi2.sh N   (multi-GPU parity check + bench under torchrun)
mkdir -p gpurun_out
N=${1:-2}
T0=$(date +%s)
stamp() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a gpurun_out/timeline_n$N.log; }
stamp "multi_gpu_check"
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py > gpurun_out/multi_check_n$N.json 2> gpurun_out/multi_check_n$N.err
echo "rc=$?" >> gpurun_out/multi_check_n$N.err
stamp "bench"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
echo "rc=$?" >> gpurun_out/bench_n$N.err
stamp "reference arm under torchrun"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
stamp "done"
tail -n 4 gpurun_out/multi_check_n$N.err; grep "\[bench\]\|rc=\|Error\|error" gpurun_out/bench_n$N.err | tail -n 12
head -c 600 gpurun_out/bench_n$N.json
