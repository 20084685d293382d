# thermo leg vs batch size: is the 2^22-cell leading dimension (2^25-byte stride between species rows) aliasing memory channels?
for c in 4194304 4000000 4194560 3000000 8388608; do
python bench.py --cells $c --steps 1 --warmup 1 --no-cpu-baseline --thermo-only 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])['thermo']; print($c, {k:(round(x['ms'],4), round(x.get('achieved_GBps',0))) for k,x in d.items()})"
done
