# round 2, GPU run 22: final validation of the committed tree -- full GPU suite, smoke, default bench line, reference arm
mkdir -p gpurun_out/r2_run22
O=gpurun_out/r2_run22
python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log; tail -2 $O/smoke.log
( time python bench.py ) > $O/bench_default.json 2> $O/bench_default.err; tail -4 $O/bench_default.err
( time python bench.py --impl reference --steps 2 --warmup 1 ) > $O/bench_reference.json 2> $O/bench_reference.err; tail -4 $O/bench_reference.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r2_run22/bench_default.json').read().strip().splitlines()[-1])
for k in ('value', 'ms_per_step', 'e2e', 'gpu_launches', 'clocks'):
    print(k, d[k])
print('roofline', {k: d['roofline'][k] for k in ('kernel', 'frac', 'achieved', 'peak')})
print('configs', {k: v.get('value') for k, v in d['configs'].items()})
print('flow', {k: v['mean_attempts_per_cell'] for k, v in d['flow_steps'].items() if isinstance(v, dict) and 'mean_attempts_per_cell' in v})
print('cpu', d['cpu_baseline'])
r = json.loads(open('gpurun_out/r2_run22/bench_reference.json').read().strip().splitlines()[-1])
print('reference arm', r['value'], r['cpu_baseline'])
PY
