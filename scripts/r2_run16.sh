# round 2, GPU run 16: warps per CTA of the warp-per-cell kernel for mechanisms with <= 24 active species (drm19, C5)
mkdir -p gpurun_out/r2_run16
O=gpurun_out/r2_run16
LIB=deepflame-dev_b200/libdfchem_b200.so
B="--cells 1048576 --steps 3 --warmup 2 --no-thermo --no-cpu-baseline --no-parity --no-strong --no-configs --no-flow"
for v in s16 s12 s10 s8; do
  cp deepflame-dev_b200/variants/lib_$v.so $LIB; touch $LIB
  python bench.py --mech drm19 $B > $O/bench_drm19_$v.json 2> $O/bench_drm19_$v.err
  DFCHEM_SOLVER=warp python bench.py --mech h2_li $B > $O/bench_h2_li_$v.json 2> $O/bench_h2_li_$v.err
done
for f in $O/bench_*.json; do echo $f; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(' value %.4g ms %.2f frac %s attempts %.2f' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['integrator_stats']['mean_attempts_per_cell']))
except Exception as e:
    print(' unreadable', e); print(open(sys.argv[1].replace('.json', '.err')).read()[-600:])
PY
done
