# round 2, GPU run 20: throughput of solve_kernel<34,2> (gri30) against the number of resident warps per SM (DFCHEM_MAX_WARPS)
mkdir -p gpurun_out/r2_run20
O=gpurun_out/r2_run20
B="--cells 262144 --steps 2 --warmup 1 --no-thermo --no-cpu-baseline --no-parity --no-strong --no-configs --no-flow"
for w in 1 2 4 6 8; do
  DFCHEM_MAX_WARPS=$w python bench.py --mech gri30 $B > $O/bench_gri30_w$w.json 2> $O/bench_gri30_w$w.err
  DFCHEM_MAX_WARPS=$w python bench.py --mech drm19 $B > $O/bench_drm19_w$w.json 2> $O/bench_drm19_w$w.err
done
for w in 12 16; do
  DFCHEM_MAX_WARPS=$w python bench.py --mech drm19 $B > $O/bench_drm19_w$w.json 2> $O/bench_drm19_w$w.err
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
