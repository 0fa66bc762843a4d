#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_expr.py -x -q > gpurun_out/pytest_expr.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_expr.log
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu.log
{ python tools/prof_mpc.py 65536 50 30; python tools/prof_mpc.py 65536 50 30 rng; python tools/prof_mpc.py 524288 100 10 rng; python tools/prof_mpc.py 65536 50 30 rng f32; } > gpurun_out/prof_mpc2.log 2>&1
