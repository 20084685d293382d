# round 2, GPU run 21: thermo_kernel with the deferred second pass (cells whose root crosses a switch temperature are queued)
mkdir -p gpurun_out/r2_run21
O=gpurun_out/r2_run21
python -m pytest tests -m gpu -q -k "thermo or c5 or transport or patch" > $O/pytest_thermo.log 2>&1; echo "pytest rc=$?" >> $O/pytest_thermo.log
tail -3 $O/pytest_thermo.log
for mech in gri30 drm19 es80; do
  python bench.py --mech $mech --cells 4194304 --thermo-only --no-cpu-baseline > $O/thermo_$mech.json 2> $O/thermo_$mech.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2_run21/thermo_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])['thermo']
        print(f.split('/')[-1], {k: (round(v['ms'], 4), round(v.get('achieved_GBps', 0))) for k, v in d.items() if 'mu' not in k})
    except Exception as e:
        print(f, 'unreadable', e)
PY
