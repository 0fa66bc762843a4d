mkdir -p gpurun_out
for r in 1 2; do
for f in ab/0_v8.so catenary-model-estimation-and-mpc-control-for-rov-tethered-systems_b200/libcatenary_b200.so; do
  echo "== $f"
  CAT_B200_LIB=$PWD/$f python bench.py --steps 50 --warmup 5 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); e=d['e2e']; print('e2e ms', e['ms_per_step'], 'rng ms', e['device_sampled_controls']['ms_per_step'], 'value', d['value'], 'p50', d['p50_ms'])"
done
done
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
