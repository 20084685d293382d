# round 2, GPU run 23: small batches -- lane-per-cell against warp-per-cell kernel by batch size (per-call policy threshold)
mkdir -p gpurun_out/r2_run23
O=gpurun_out/r2_run23
for mech in es80 h2_li c7_onestep; do
 for cells in 1000 2048 4096 8192 16384 65536; do
  B="--mech $mech --cells $cells --steps 5 --warmup 3 --no-thermo --no-cpu-baseline --no-parity --no-strong --no-flow"
  DFCHEM_SOLVER=lane python bench.py $B > $O/b_${mech}_${cells}_lane.json 2> $O/err
  DFCHEM_SOLVER=warp python bench.py $B > $O/b_${mech}_${cells}_warp.json 2> $O/err
 done
done
python - <<'PY'
import json, glob, re
rows = {}
for f in sorted(glob.glob('gpurun_out/r2_run23/b_*.json')):
    m = re.match(r'.*/b_(.+)_(\d+)_(lane|warp)\.json', f)
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        rows.setdefault((m.group(1), int(m.group(2))), {})[m.group(3)] = d['ms_per_step']
    except Exception as e:
        print(f, 'unreadable', e)
for k in sorted(rows):
    print(k, {a: round(b, 3) for a, b in rows[k].items()})
PY
