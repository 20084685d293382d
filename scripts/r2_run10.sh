# round 2, GPU run 10: lane-per-cell kernel variants on the small mechanisms
#   v0 = committed gather form (CSR lists, unrolled stages)      v1 = ELL lists + rolled stage loop
#   v2 = v1 + reaction loop by pairs                             v3 = v2 + LU factors kept in registers (NAP <= 6)
# then the lane-forced parity tests on v3 and the es80 default-policy check
mkdir -p gpurun_out/r2_run10
O=gpurun_out/r2_run10
LIB=deepflame-dev_b200/libdfchem_b200.so
B="--cells 1048576 --steps 3 --warmup 3 --no-thermo --no-cpu-baseline --no-parity --no-strong"
for v in v0 v1 v2 v3; do
  cp deepflame-dev_b200/variants/lib_$v.so $LIB; touch $LIB
  for mech in es80 h2_li c7_onestep; do
    DFCHEM_SOLVER=lane python bench.py --mech $mech $B > $O/bench_${mech}_$v.json 2> $O/bench_${mech}_$v.err
  done
done
python bench.py --mech es80 $B > $O/bench_es80_v3_default.json 2> $O/bench_es80_v3_default.err
DFCHEM_SOLVER=warp python bench.py --mech h2_li $B > $O/bench_h2_li_warp.json 2> $O/bench_h2_li_warp.err
DFCHEM_SOLVER=lane python -m pytest tests/test_gpu_parity.py tests/test_equal_tolerance.py tests/test_mock_caller.py -m gpu -q -k "es80 or h2_li or c7" > $O/pytest_lane.log 2>&1; echo "pytest(lane) rc=$?" >> $O/pytest_lane.log
tail -4 $O/pytest_lane.log
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(' value %.4g ms %.2f frac %s attempts %.2f e2e %.4g' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['integrator_stats']['mean_attempts_per_cell'], d['e2e']['value']))
except Exception as e:
    print(' unreadable', e); print(open(sys.argv[1].replace('.json', '.err')).read()[-600:])
PY
done
