# thermo variant e + transport CTAs/SM sweep (thermo-only bench legs), then GPU tests
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
for c in 3 4 5 6 8; do
DFCHEM_TR_CTAS=$c python bench.py --cells 4194304 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only > gpurun_out/thermo_ctas$c.json 2> gpurun_out/thermo_ctas$c.err
done
python bench.py --mech es80 --cells 1048576 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only > gpurun_out/thermo_es80.json 2>&1
python bench.py --mech drm19 --cells 1048576 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only > gpurun_out/thermo_drm19.json 2>&1
BT="python bench.py --cells 4194304 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only"
ncu --set full --clock-control none --import-source on -k regex:thermo_kernel -s 6 -c 2 -o gpurun_out/prof_thermo_e $BT > gpurun_out/ncu3e.log 2>&1
tail -3 gpurun_out/pytest_gpu.log; for c in 3 4 5 6 8; do python -c "
import json,sys
d=json.load(open('gpurun_out/thermo_ctas$c.json'))['thermo']
print($c, {k:(round(v['ms'],3)) for k,v in d.items()})"; done
