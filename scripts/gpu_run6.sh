# solve kernel: reciprocal pivot computed during the previous Gauss-Jordan step (new) vs base; transport unroll/min-CTA variants
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu6.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu6.log
B="python bench.py --cells 1048576 --steps 3 --warmup 2 --no-cpu-baseline --no-thermo"
for v in base new base new; do
L=deepflame-dev_b200/libdfchem_b200_base.so; [ $v = new ] && L=deepflame-dev_b200/libdfchem_b200.so
DFCHEM_LIB=$PWD/$L $B 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('solve $v', d['value'], d['ms_per_step'], d['integrator_stats']['accepted_steps'], d['integrator_stats']['rejected_steps'])"
done
for v in tr88 tr86 tr66; do
DFCHEM_LIB=$PWD/deepflame-dev_b200/libdfchem_b200_$v.so python bench.py --cells 4194304 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])['thermo']; print('$v', {k:round(x['ms'],3) for k,x in d.items()})"
done
tail -2 gpurun_out/pytest_gpu6.log
