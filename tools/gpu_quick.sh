#!/bin/bash
# quick single-GPU pass: parity tests + default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py > gpurun_out/bench_quick.json 2> gpurun_out/bench_quick.err
