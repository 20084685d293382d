# round 2, GPU run 5: two-cells-per-thread thermo kernel, register budget / batch sweep (gri30 4M, drm19 4M, es80 4M cells)
mkdir -p gpurun_out/r2_run5
O=gpurun_out/r2_run5
python -m pytest tests -m gpu -q -k "thermo or c5" > $O/pytest_thermo.log 2>&1; echo "pytest rc=$?" >> $O/pytest_thermo.log
tail -3 $O/pytest_thermo.log
for cfg in 26 28 34 36 43 44; do
  DFCHEM_THERMO_CFG=$cfg python bench.py --mech gri30 --thermo-only --no-cpu-baseline > $O/thermo_gri30_pair$cfg.json 2> $O/thermo_gri30_pair$cfg.err
done
DFCHEM_THERMO=ldg python bench.py --mech gri30 --thermo-only --no-cpu-baseline > $O/thermo_gri30_ldg.json 2> $O/err
for mech in drm19 es80; do
  python bench.py --mech $mech --cells 4194304 --thermo-only --no-cpu-baseline > $O/thermo_${mech}_pair36.json 2> $O/err
  DFCHEM_THERMO_CFG=26 python bench.py --mech $mech --cells 4194304 --thermo-only --no-cpu-baseline > $O/thermo_${mech}_pair26.json 2> $O/err
  DFCHEM_THERMO=ldg python bench.py --mech $mech --cells 4194304 --thermo-only --no-cpu-baseline > $O/thermo_${mech}_ldg.json 2> $O/err
done
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r2_run5/thermo_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])['thermo']
        print(f.split('/')[-1], {k: (round(v['ms'], 4), round(v.get('achieved_GBps', 0))) for k, v in d.items() if 'mu' not in k})
    except Exception as e:
        print(f, 'unreadable', e)
PY
