# round 2, GPU run 13 (8 GPUs): bench at N = 8 and N = 4 (weak-scaled main leg + strong-scaled skewed leg with the
# in-library NCCL rebalancer, allAverage and headTail), and the 2-rank NCCL balance tests
mkdir -p gpurun_out/r2_run13
O=gpurun_out/r2_run13
python -m pytest tests/test_gpu_balance.py -m gpu -q > $O/pytest_balance.log 2>&1; echo "pytest rc=$?" >> $O/pytest_balance.log
tail -3 $O/pytest_balance.log
T="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
A="--steps 2 --warmup 1 --no-cpu-baseline --no-thermo --no-parity"
timeout 600 $T --nproc-per-node 8 --master-port 29611 bench.py --gpus 8 $A > $O/bench_n8.json 2> $O/bench_n8.err; echo "n8 rc=$?"
timeout 600 $T --nproc-per-node 8 --master-port 29612 bench.py --gpus 8 $A --balance-algorithm headTail --steps 1 > $O/bench_n8_headtail.json 2> $O/bench_n8_headtail.err; echo "n8 headTail rc=$?"
timeout 600 $T --nproc-per-node 4 --master-port 29613 bench.py --gpus 4 $A > $O/bench_n4.json 2> $O/bench_n4.err; echo "n4 rc=$?"
for f in $O/bench_n*.json; do echo $f; python - "$f" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(' value %.4g ms %.1f n_gpus %d' % (d['value'], d['ms_per_step'], d['n_gpus']))
    print(' strong', json.dumps(d.get('strong'))[:1200])
except Exception as e:
    print(' unreadable', e); print(open(sys.argv[1].replace('.json', '.err')).read()[-800:])
PY
done
