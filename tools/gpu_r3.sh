#!/bin/bash
mkdir -p gpurun_out
{
echo "default(minb5)"; python tools/prof_shape.py; python tools/prof_shape.py 1048576 15; python tools/prof_shape.py 1048576 8 f32
for mb in 6 7 8; do echo "minb $mb"; CAT_B200_LIB=$PWD/tools/variants/lib_minb$mb.so python tools/prof_shape.py; CAT_B200_LIB=$PWD/tools/variants/lib_minb$mb.so python tools/prof_shape.py 1048576 15; done
} > gpurun_out/shape_variants.log 2>&1
python -m pytest tests -m gpu -x -q -k "shape or transform or catenary_class" > gpurun_out/pytest_shape.log 2>&1
