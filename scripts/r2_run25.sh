# round 2, GPU run 25: lane-per-cell kernel, refill granularity (idle lanes that trigger a fetch from the queue): 2 / 4 / 8 / 16
mkdir -p gpurun_out/r2_run25
O=gpurun_out/r2_run25
LIB=deepflame-dev_b200/libdfchem_b200.so
B="--cells 1048576 --steps 3 --warmup 3 --no-thermo --no-cpu-baseline --no-parity --no-strong --no-flow"
for v in r2 r4 r8 r16; do
  cp deepflame-dev_b200/variants/lib_$v.so $LIB; touch $LIB
  for mech in es80 h2_li c7_onestep; do
    python bench.py --mech $mech $B > $O/bench_${mech}_$v.json 2> $O/bench_${mech}_$v.err
  done
done
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(' value %.4g ms %.2f' % (d['value'], d['ms_per_step']))
except Exception as e:
    print(' unreadable', e); print(open(sys.argv[1].replace('.json', '.err')).read()[-600:])
PY
done
