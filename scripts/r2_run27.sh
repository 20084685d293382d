# round 2, GPU run 27 (2 GPUs): NCCL balance tests and the N = 2 bench line on the final tree
mkdir -p gpurun_out/r2_run27
O=gpurun_out/r2_run27
python -m pytest tests/test_gpu_balance.py -m gpu -q > $O/pytest_balance.log 2>&1; echo "pytest rc=$?" >> $O/pytest_balance.log
tail -3 $O/pytest_balance.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 2 --warmup 1 --no-cpu-baseline --no-thermo --no-parity > $O/bench_n2.json 2> $O/bench_n2.err; echo "n2 rc=$?"
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r2_run27/bench_n2.json').read().strip().splitlines()[-1])
print('value %.4g n_gpus %d' % (d['value'], d['n_gpus'])); s = d['strong']
print({k: s[k] for k in ('load_imbalance_max_over_mean', 'speedup', 'results_bit_identical_to_unbalanced', 'exchange_fraction_of_step')}, s['without_rebalance']['value'], s['with_rebalance']['value'])
PY
