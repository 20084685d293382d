# round 2, GPU run 15: consecutive-flow-steps leg (deltaTChem carried vs cold start), gri30 / drm19 / es80
mkdir -p gpurun_out/r2_run15
O=gpurun_out/r2_run15
B="--cells 1048576 --steps 1 --warmup 1 --no-thermo --no-cpu-baseline --no-parity --no-configs --no-strong"
for mech in gri30 drm19 es80; do
  timeout 600 python bench.py --mech $mech $B > $O/bench_flow_$mech.json 2> $O/bench_flow_$mech.err; echo "$mech rc=$?"
  python - "$O/bench_flow_$mech.json" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(json.dumps(d.get('flow_steps'), indent=0)[:2500])
except Exception as e:
    print(' unreadable', e); print(open(sys.argv[1].replace('.json', '.err')).read()[-1500:])
PY
done
