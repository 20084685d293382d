# round 2, GPU run 26: lane kernel with run-time refill granularity (4; 16 for one-step mechanisms): benches + lane parity tests
mkdir -p gpurun_out/r2_run26
O=gpurun_out/r2_run26
B="--cells 1048576 --steps 3 --warmup 3 --no-thermo --no-cpu-baseline --no-parity --no-strong --no-flow"
for mech in es80 h2_li c7_onestep; do
  python bench.py --mech $mech $B > $O/bench_${mech}.json 2> $O/bench_${mech}.err
done
python -m pytest tests/test_gpu_parity.py tests/test_equal_tolerance.py tests/test_mock_caller.py tests/test_gpu_properties.py -m gpu -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -4 $O/pytest.log
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(' value %.4g ms %.2f e2e %.4g kernel %s' % (d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['kernel']))
except Exception as e:
    print(' unreadable', e); print(open(sys.argv[1].replace('.json', '.err')).read()[-600:])
PY
done
