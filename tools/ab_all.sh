#!/bin/bash
# A/B of library variants in ab/*.so on one box: MPC rollout (explicit, sampled), shape kernel (final / original /
# all four curves, fp64 and fp32), then the GPU parity tests with the in-tree library.  usage: tools/ab_all.sh [rounds]
R=${1:-2}
mkdir -p gpurun_out
for r in $(seq $R); do
  for f in ab/*.so; do
    echo "== $f"
    CAT_B200_LIB=$PWD/$f python tools/prof_mpc.py 65536 50 60 | cut -d'(' -f1
    CAT_B200_LIB=$PWD/$f python tools/prof_mpc.py 65536 50 60 rng | cut -d'(' -f1
    CAT_B200_LIB=$PWD/$f python tools/prof_shape.py 1048576 8 f64 20
    CAT_B200_LIB=$PWD/$f python tools/prof_shape.py 1048576 1 f64 20
    CAT_B200_LIB=$PWD/$f python tools/prof_shape.py 1048576 15 f64 10
    CAT_B200_LIB=$PWD/$f python tools/prof_shape.py 1048576 8 f32 20
  done
done
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
