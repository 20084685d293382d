# thermo variant f: register budget sweep (min CTAs/SM 4/3/2 -> 64/80/126 registers), thermo-only bench legs
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "thermo or transport or edge" > gpurun_out/pytest_gpu4.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu4.log
for v in 4 3 2; do
L=deepflame-dev_b200/libdfchem_b200_minb$v.so; [ $v = 4 ] && L=deepflame-dev_b200/libdfchem_b200.so
DFCHEM_LIB=$PWD/$L python bench.py --cells 4194304 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only > gpurun_out/thermo_minb$v.json 2> gpurun_out/thermo_minb$v.err
DFCHEM_LIB=$PWD/$L python bench.py --mech es80 --cells 1048576 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only > gpurun_out/thermo_es80_minb$v.json 2> gpurun_out/thermo_es80_minb$v.err
done
tail -3 gpurun_out/pytest_gpu4.log; for v in 4 3 2; do python -c "
import json,sys
for f in ('thermo_minb$v','thermo_es80_minb$v'):
    d=json.load(open('gpurun_out/'+f+'.json'))['thermo']
    print(f, {k:(round(v['ms'],4)) for k,v in d.items()})"; done
