#!/bin/bash
# A/B timing of library variants in ab/*.so on one box (box-to-box clocks differ by a few %): alternates the
# variants so that drift hits all of them alike.  usage: tools/ab.sh [rounds]
R=${1:-3}
for r in $(seq $R); do
  for f in ab/*.so; do
    echo "== $f"
    CAT_B200_LIB=$PWD/$f python tools/prof_mpc.py 65536 50 60 | cut -d'(' -f1
    CAT_B200_LIB=$PWD/$f python tools/prof_mpc.py 65536 50 60 rng | cut -d'(' -f1
    CAT_B200_LIB=$PWD/$f python tools/prof_mpc.py 65536 50 60 x f32 | cut -d'(' -f1
    if [[ "${AB_SHAPES:-0}" == 1 ]]; then
      CAT_B200_LIB=$PWD/$f python tools/prof_shape.py 1048576 8 f64 20
      CAT_B200_LIB=$PWD/$f python tools/prof_shape.py 1048576 8 f32 20
      CAT_B200_LIB=$PWD/$f python tools/prof_shape.py 1048576 15 f64 10
      CAT_B200_LIB=$PWD/$f python tools/prof_shape.py 1048576 1 f64 20
    fi
    CAT_B200_LIB=$PWD/$f python tools/prof_loop.py
  done
done
