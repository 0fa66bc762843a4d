#!/bin/bash
mkdir -p gpurun_out
{ python tools/prof_mpc.py 65536 50 40; python tools/prof_mpc.py 65536 50 40 rng; python tools/prof_mpc.py 56832 50 40; python tools/prof_mpc.py 75776 50 40; python tools/prof_mpc.py 65536 50 40 x f32; } > gpurun_out/prof_mpc3.log 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q -k "mpc or loop or controller or smoke or device_step" > gpurun_out/pytest_mpc.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_mpc.log
