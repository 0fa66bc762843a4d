#!/bin/bash
mkdir -p gpurun_out
{ python tools/prof_shape.py 1048576 8 f64 30; python tools/prof_shape.py 1048576 15 f64 30; python tools/prof_shape.py 1048576 15 f64 30; python tools/prof_shape.py 1048576 1 f64 30; } > gpurun_out/shape3.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shape -s 1 -c 1 -f -o gpurun_out/shape_full2 python tools/prof_shape.py > gpurun_out/ncu_shape2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_shape -s 1 -c 1 -f -o gpurun_out/shape_full2_w15 python tools/prof_shape.py 1048576 15 > gpurun_out/ncu_shape2b.log 2>&1
