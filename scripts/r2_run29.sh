# round 2, GPU run 29 (8 GPUs): the bench line exactly as the driver launches it at N = 8 (default legs)
mkdir -p gpurun_out/r2_run29
O=gpurun_out/r2_run29
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29911 bench.py --gpus 8 --steps 3 --warmup 3 ) > $O/bench_n8_default.json 2> $O/bench_n8_default.err; echo "rc=$?"
tail -5 $O/bench_n8_default.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r2_run29/bench_n8_default.json').read().strip().splitlines()[-1])
print('value %.4g n_gpus %d e2e %.4g' % (d['value'], d['n_gpus'], d['e2e']['value']), d['clocks'])
s = d['strong']; print({k: s[k] for k in ('load_imbalance_max_over_mean', 'speedup', 'results_bit_identical_to_unbalanced')}, s['with_rebalance']['value'])
print({k: (v.get('ms'), v.get('cells_on_a_nasa_switch_temperature')) for k, v in d['thermo'].items()})
PY
