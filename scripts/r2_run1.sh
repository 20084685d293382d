# round 2, GPU run 1: parity tests, lane-per-cell vs warp-per-cell on the small mechanisms, gri30 line with the new controller
mkdir -p gpurun_out/r2_run1
O=gpurun_out/r2_run1
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
for mech in es80 h2_li c7_onestep; do
  python bench.py --mech $mech --cells 1048576 --steps 3 --warmup 3 --no-thermo > $O/bench_${mech}_lpc.json 2> $O/bench_${mech}_lpc.err
  DFCHEM_SOLVER=warp python bench.py --mech $mech --cells 1048576 --steps 3 --warmup 3 --no-thermo --no-cpu-baseline > $O/bench_${mech}_warp.json 2> $O/bench_${mech}_warp.err
done
python bench.py --steps 2 --warmup 3 > $O/bench_gri30.json 2> $O/bench_gri30.err
tail -3 $O/pytest_gpu.log
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(' value %.4g ms %.2f frac %s attempts %.2f failed %s e2e %.4g' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['integrator_stats']['mean_attempts_per_cell'], d['integrator_stats']['failed_cells'], d['e2e']['value']))
except Exception as e:
    print(' unreadable', e)
PY
done
