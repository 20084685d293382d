# round 2, GPU run 4: bulk-async thermo kernel vs the register-staged one (gri30 4M cells, drm19, es80), thermo tests
mkdir -p gpurun_out/r2_run4
O=gpurun_out/r2_run4
python -m pytest tests -m gpu -q -k "thermo or transport or c5" > $O/pytest_thermo.log 2>&1; echo "pytest rc=$?" >> $O/pytest_thermo.log
tail -4 $O/pytest_thermo.log
for mech in gri30 drm19 es80; do
  python bench.py --mech $mech --thermo-only --no-cpu-baseline > $O/thermo_${mech}_tma.json 2> $O/thermo_${mech}_tma.err
  DFCHEM_THERMO=ldg python bench.py --mech $mech --thermo-only --no-cpu-baseline > $O/thermo_${mech}_ldg.json 2> $O/thermo_${mech}_ldg.err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2_run4/thermo_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])['thermo']
        print(f.split('/')[-1], {k: (round(v['ms'], 4), round(v.get('achieved_GBps', 0))) for k, v in d.items()})
    except Exception as e:
        print(f, 'unreadable', e)
PY
tail -3 $O/thermo_gri30_tma.err
