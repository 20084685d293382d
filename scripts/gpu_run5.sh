# thermo variant g (ping-pong pipeline): register budget x waves sweep
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "thermo or transport or edge" > gpurun_out/pytest_gpu5.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu5.log
for v in 4 3; do for w in 1 2; do
L=deepflame-dev_b200/libdfchem_b200_minb$v.so; [ $v = 4 ] && L=deepflame-dev_b200/libdfchem_b200.so
DFCHEM_THERMO_WAVES=$w DFCHEM_LIB=$PWD/$L python bench.py --cells 4194304 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only > gpurun_out/thermo_g_minb${v}_w$w.json 2> gpurun_out/thermo_g.err
done; done
tail -3 gpurun_out/pytest_gpu5.log; for v in 4 3; do for w in 1 2; do python -c "
import json,sys
f='thermo_g_minb${v}_w$w'
d=json.load(open('gpurun_out/'+f+'.json'))['thermo']
print(f, {k:(round(v['ms'],4)) for k,v in d.items()})"; done; done
