# A/B of k_shape library variants in ab/*.so (tools/prof_shape.py: 1,048,576 configs x 100 points;
# mask bits: 1 original, 2 theta-rotated, 4 theta-aligned, 8 final)
for r in 1 2; do
  for f in ab/*.so; do
    echo "== $f"
    for dt in f64 f32; do
      for w in 8 1 15 9; do
        CAT_B200_LIB=$PWD/$f python tools/prof_shape.py 1048576 $w $dt 15
      done
    done
  done
done
