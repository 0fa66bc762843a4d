#!/bin/bash
# compute-sanitizer passes over small launches of the two hot kernels and the controller (memcheck, racecheck,
# initcheck-free: torch's caching allocator hands out uninitialised blocks by design).  usage: gpurun -- bash tools/sanitize.sh
mkdir -p gpurun_out
S=/usr/local/cuda/bin/compute-sanitizer
O=gpurun_out/sanitize.log
: > $O
for tool in memcheck racecheck synccheck; do
  echo "=== $tool prof_mpc explicit" >> $O
  timeout 600 $S --tool $tool --error-exitcode 9 python tools/prof_mpc.py 8192 20 2 >> $O 2>&1; echo "rc=$?" >> $O
  echo "=== $tool prof_mpc rng" >> $O
  timeout 600 $S --tool $tool --error-exitcode 9 python tools/prof_mpc.py 8192 20 2 rng >> $O 2>&1; echo "rc=$?" >> $O
  echo "=== $tool prof_shape final" >> $O
  timeout 600 $S --tool $tool --error-exitcode 9 python tools/prof_shape.py 20000 8 f64 1 >> $O 2>&1; echo "rc=$?" >> $O
  echo "=== $tool prof_shape all four" >> $O
  timeout 600 $S --tool $tool --error-exitcode 9 python tools/prof_shape.py 20000 15 f64 1 >> $O 2>&1; echo "rc=$?" >> $O
done
echo "=== memcheck smoke" >> $O
timeout 900 $S --tool memcheck --error-exitcode 9 python -c "import __graft_entry__ as g; g.smoke()" >> $O 2>&1; echo "rc=$?" >> $O
grep -E "^===|rc=|ERROR SUMMARY|RACECHECK SUMMARY" $O
