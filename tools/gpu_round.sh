#!/bin/bash
# One GPU-box pass: parity tests, bench (ours + reference arm), ncu launch list and full captures.
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1
python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
python tools/prof_mpc.py 65536 50 20 > gpurun_out/prof_mpc.log 2>&1
python tools/prof_mpc.py 65536 50 20 rng >> gpurun_out/prof_mpc.log 2>&1
python tools/prof_mpc.py 65536 50 20 x f32 >> gpurun_out/prof_mpc.log 2>&1
python tools/prof_shape.py > gpurun_out/prof_shape.log 2>&1
python tools/prof_shape.py 1048576 1 >> gpurun_out/prof_shape.log 2>&1
python tools/prof_shape.py 1048576 15 >> gpurun_out/prof_shape.log 2>&1
python tools/prof_shape.py 1048576 8 f32 >> gpurun_out/prof_shape.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 20 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_mpc_rollout -s 3 -c 2 -f -o gpurun_out/mpc_full python tools/prof_mpc.py 65536 50 8 > gpurun_out/ncu_mpc.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shape -s 1 -c 1 -f -o gpurun_out/shape_full python tools/prof_shape.py > gpurun_out/ncu_shape.log 2>&1
ls -la gpurun_out
ncu --set full --clock-control none --import-source on -k regex:k_mpc_rollout -s 3 -c 1 -f -o gpurun_out/mpc_f32 python tools/prof_mpc.py 65536 50 8 x f32 > gpurun_out/ncu_f32.log 2>&1
