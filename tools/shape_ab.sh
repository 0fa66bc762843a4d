# A/B of k_shape's two-stores-per-point form (CAT_SHAPE_ST2=1) against three scalar stores, by output mask
# (tools/prof_shape.py: 1,048,576 configs x 100 points; mask bits: 1 original, 2 theta-rotated, 4 theta-aligned, 8 final)
for r in 1 2; do
  for st2 in 0 1; do
    echo "== CAT_SHAPE_ST2=$st2"
    for dt in f64 f32; do
      for w in 3 9 12 14 15; do
        CAT_SHAPE_ST2=$st2 python tools/prof_shape.py 1048576 $w $dt 10
      done
    done
    CAT_SHAPE_ST2=$st2 python tools/prof_shape.py 1048576 8 f32 20
    CAT_SHAPE_ST2=$st2 python tools/prof_shape.py 1048576 1 f32 20
  done
done
