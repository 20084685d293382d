# round 2, GPU run 11: final lane kernel (ELL + rolled + LU in registers, no LU region) on the small mechanisms, then one
# full ncu capture of solve_kernel<34,2> (gri30) with source import for the per-line stall table
mkdir -p gpurun_out/r2_run11
O=gpurun_out/r2_run11
B="--cells 1048576 --steps 3 --warmup 3 --no-thermo --no-cpu-baseline --no-parity --no-strong"
for mech in es80 h2_li c7_onestep; do
  python bench.py --mech $mech $B > $O/bench_${mech}_v4.json 2> $O/bench_${mech}_v4.err
done
python -m pytest tests/test_gpu_parity.py tests/test_equal_tolerance.py tests/test_mock_caller.py tests/test_gpu_properties.py -m gpu -q > $O/pytest_default.log 2>&1; echo "pytest rc=$?" >> $O/pytest_default.log
tail -4 $O/pytest_default.log
G="python bench.py --cells 131072 --steps 1 --warmup 1 --no-cpu-baseline --no-configs --no-parity --no-thermo --no-strong"
$G > $O/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:solve_kernel -c 1 -o $O/prof_solve $G > $O/ncu.log 2>&1
tail -3 $O/ncu.log
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(' value %.4g ms %.2f frac %s attempts %.2f e2e %.4g' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['integrator_stats']['mean_attempts_per_cell'], d['e2e']['value']))
except Exception as e:
    print(' unreadable', e); print(open(sys.argv[1].replace('.json', '.err')).read()[-600:])
PY
done
ls -la $O
