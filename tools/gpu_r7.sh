#!/bin/bash
mkdir -p gpurun_out
{ python tools/prof_shape.py; python tools/prof_shape.py 1048576 15; python tools/prof_shape.py 1048576 1; python tools/prof_shape.py 1048576 8 f32; } > gpurun_out/shape2.log 2>&1
python -m pytest tests -m gpu -x -q -k "shape or transform or catenary_class or smoke" > gpurun_out/pytest_shape.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_shape.log
