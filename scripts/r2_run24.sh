# round 2, GPU run 24: per-call kernel policy (small batches one warp per cell) -- full GPU suite, 1000-cell es80 sub-step by default
mkdir -p gpurun_out/r2_run24
O=gpurun_out/r2_run24
python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -5 $O/pytest_gpu.log
for cells in 1000 1048576; do
  python bench.py --mech es80 --cells $cells --steps 5 --warmup 3 --no-thermo --no-cpu-baseline --no-parity --no-strong --no-flow > $O/b_es80_$cells.json 2> $O/err
  python - "$O/b_es80_$cells.json" <<'PY'
import json, sys
d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], 'ms %.3f value %.4g kernel %s' % (d['ms_per_step'], d['value'], d['roofline']['kernel']))
PY
done
