# round 2, GPU run 7: gather-form lane-per-cell kernel (lpc v2) vs warp-per-cell on the small mechanisms; lane-forced parity tests
mkdir -p gpurun_out/r2_run7
O=gpurun_out/r2_run7
DFCHEM_SOLVER=lane python -m pytest tests/test_gpu_parity.py tests/test_equal_tolerance.py tests/test_mock_caller.py -m gpu -q -k "es80 or h2_li or c7" > $O/pytest_lane.log 2>&1; echo "pytest(lane) rc=$?" >> $O/pytest_lane.log
tail -4 $O/pytest_lane.log
for mech in es80 h2_li c7_onestep; do
  DFCHEM_SOLVER=lane python bench.py --mech $mech --cells 1048576 --steps 3 --warmup 3 --no-thermo --no-cpu-baseline > $O/bench_${mech}_lane.json 2> $O/bench_${mech}_lane.err
  DFCHEM_SOLVER=warp python bench.py --mech $mech --cells 1048576 --steps 3 --warmup 3 --no-thermo --no-cpu-baseline > $O/bench_${mech}_warp.json 2> $O/bench_${mech}_warp.err
done
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(' value %.4g ms %.2f frac %s attempts %.2f e2e %.4g' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['integrator_stats']['mean_attempts_per_cell'], d['e2e']['value']))
except Exception as e:
    print(' unreadable', e); print(open(sys.argv[1].replace('.json', '.err')).read()[-600:])
PY
done
