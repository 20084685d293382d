# round 2, GPU run 12: warp-per-cell kernel variants on gri30 / drm19 (loads-before-stores restructuring)
#   w0 = committed   w1 = chunked run_wdot + reaction pairs   w2 = chunked run_wdot only
#   w3 = w1 + Jacobian row program by pairs   w4 = row pairs only
mkdir -p gpurun_out/r2_run12
O=gpurun_out/r2_run12
LIB=deepflame-dev_b200/libdfchem_b200.so
B="--cells 1048576 --steps 2 --warmup 1 --no-thermo --no-cpu-baseline --no-parity --no-strong --no-configs"
for v in w0 w1 w2 w3 w4; do
  cp deepflame-dev_b200/variants/lib_$v.so $LIB; touch $LIB
  for mech in gri30 drm19; do
    python bench.py --mech $mech $B > $O/bench_${mech}_$v.json 2> $O/bench_${mech}_$v.err
  done
done
cp deepflame-dev_b200/variants/lib_w3.so $LIB; touch $LIB
python -m pytest tests/test_gpu_parity.py tests/test_equal_tolerance.py tests/test_gpu_properties.py tests/test_mechanism_classes.py -m gpu -q > $O/pytest_w3.log 2>&1; echo "pytest rc=$?" >> $O/pytest_w3.log
tail -4 $O/pytest_w3.log
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(' value %.4g ms %.2f frac %s attempts %.2f e2e %.4g' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['integrator_stats']['mean_attempts_per_cell'], d['e2e']['value']))
except Exception as e:
    print(' unreadable', e); print(open(sys.argv[1].replace('.json', '.err')).read()[-600:])
PY
done
