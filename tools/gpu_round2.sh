#!/bin/bash
# One GPU-box pass of round 2: parity tests, smoke, bench (ours + reference arm), ncu launch list and full captures.
# usage: gpurun --timeout 1500 -- bash tools/gpu_round2.sh [tests] [bench] [ncu] [ref]   (default: all)
mkdir -p gpurun_out
PARTS="${*:-tests bench ncu ref}"
T0=$(date +%s)
stamp() { echo "[$(( $(date +%s) - T0 )) s] $*" | tee -a gpurun_out/timeline.log; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.log 2>&1
if [[ $PARTS == *tests* ]]; then
  stamp "pytest -m gpu"
  timeout 1200 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
  tail -n 5 gpurun_out/pytest_gpu.log
  stamp "smoke"
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
  tail -n 2 gpurun_out/smoke.log
fi
if [[ $PARTS == *bench* ]]; then
  stamp "bench"
  timeout 900 python bench.py > gpurun_out/bench1.json 2> gpurun_out/bench1.err; echo "bench rc=$?" >> gpurun_out/bench1.err
  tail -n 12 gpurun_out/bench1.err
fi
if [[ $PARTS == *ref* ]]; then
  stamp "reference arm"
  timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
fi
if [[ $PARTS == *ab* ]]; then
  stamp "A/B of the library variants in ab/"
  bash tools/ab.sh 2 > gpurun_out/ab.log 2>&1
  cat gpurun_out/ab.log
fi
if [[ $PARTS == *prof* || $PARTS == *ncu* ]]; then
  stamp "prof scripts"
  python tools/prof_mpc.py 65536 50 20 > gpurun_out/prof_mpc.log 2>&1
  python tools/prof_mpc.py 65536 50 20 rng >> gpurun_out/prof_mpc.log 2>&1
  python tools/prof_mpc.py 65536 50 20 x f32 >> gpurun_out/prof_mpc.log 2>&1
  python tools/prof_shape.py > gpurun_out/prof_shape.log 2>&1
  python tools/prof_shape.py 1048576 1 >> gpurun_out/prof_shape.log 2>&1
  python tools/prof_shape.py 1048576 15 >> gpurun_out/prof_shape.log 2>&1
  python tools/prof_shape.py 1048576 8 f32 >> gpurun_out/prof_shape.log 2>&1
  cat gpurun_out/prof_mpc.log gpurun_out/prof_shape.log
fi
if [[ $PARTS == *ncu* ]]; then
  stamp "ncu launch list"
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 20 --warmup 3 > gpurun_out/ncu_launch.log 2>&1
  stamp "ncu full f64"
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_mpc_rollout -s 3 -c 2 -f -o gpurun_out/mpc_full python tools/prof_mpc.py 65536 50 8 > gpurun_out/ncu_mpc.log 2>&1
  stamp "ncu full f64 rng"
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_mpc_rollout -s 3 -c 1 -f -o gpurun_out/mpc_rng python tools/prof_mpc.py 65536 50 8 rng > gpurun_out/ncu_rng.log 2>&1
  stamp "ncu full f32"
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_mpc_rollout -s 3 -c 1 -f -o gpurun_out/mpc_f32 python tools/prof_mpc.py 65536 50 8 x f32 > gpurun_out/ncu_f32.log 2>&1
  stamp "ncu full shape"
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_shape -s 1 -c 1 -f -o gpurun_out/shape_full python tools/prof_shape.py > gpurun_out/ncu_shape.log 2>&1
fi
stamp "done"
ls -la gpurun_out | head -40
