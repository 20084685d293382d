# round 2, GPU run 28 (2 GPUs): the bench line exactly as the driver launches it at N = 2 (default legs), wall time
mkdir -p gpurun_out/r2_run28
O=gpurun_out/r2_run28
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29811 bench.py --gpus 2 --steps 3 --warmup 3 ) > $O/bench_n2_default.json 2> $O/bench_n2_default.err; echo "rc=$?"
tail -5 $O/bench_n2_default.err
( time timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29812 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 ) > $O/bench_n2_reference.json 2> $O/bench_n2_reference.err; echo "ref rc=$?"
tail -4 $O/bench_n2_reference.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r2_run28/bench_n2_default.json').read().strip().splitlines()[-1])
print(sorted(d.keys()))
print('value %.4g n_gpus %d e2e %.4g' % (d['value'], d['n_gpus'], d['e2e']['value']), d['clocks'])
r = open('gpurun_out/r2_run28/bench_n2_reference.json').read().strip().splitlines()
print(len(r), 'reference lines;', r[-1][:200])
PY
