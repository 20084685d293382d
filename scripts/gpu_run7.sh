# thermo variant e: species-batch size x register budget sweep (DFC_THERMO_BATCH, DFC_THERMO_MINB)
mkdir -p gpurun_out
for v in base t18_4 t18_3 t9_5 t36_2; do
L=deepflame-dev_b200/libdfchem_b200_$v.so; [ $v = base ] && L=deepflame-dev_b200/libdfchem_b200.so
DFCHEM_LIB=$PWD/$L python bench.py --cells 4194304 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])['thermo']; print('$v', {k:round(x['ms'],4) for k,x in d.items()})"
done
