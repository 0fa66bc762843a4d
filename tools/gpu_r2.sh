#!/bin/bash
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_controller.py -x -q > gpurun_out/pytest_ctrl.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_ctrl.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
for b in 32 64 128; do
  echo "block $b" >> gpurun_out/blocks.log
  CAT_MPC_BLOCK_SIZE=$b python tools/prof_mpc.py 65536 50 30 >> gpurun_out/blocks.log 2>&1
  CAT_MPC_BLOCK_SIZE=$b python tools/prof_mpc.py 65536 50 30 rng >> gpurun_out/blocks.log 2>&1
  CAT_MPC_BLOCK_SIZE=$b python tools/prof_mpc.py 524288 100 10 rng >> gpurun_out/blocks.log 2>&1
done
python bench.py > gpurun_out/bench2.json 2> gpurun_out/bench2.err
