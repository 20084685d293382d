# N-GPU routine (gpurun --gpus N): weak-scaling bench line, then the skewed batch with and without rebalancing
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533"
$TR bench.py --gpus $N --steps 2 --warmup 3 --no-thermo > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
$TR bench.py --gpus $N --steps 2 --warmup 3 --no-thermo --cells 1048576 --skew 4 > gpurun_out/bench_n${N}_skew.json 2> gpurun_out/bench_n${N}_skew.err
$TR bench.py --gpus $N --steps 2 --warmup 3 --no-thermo --cells 1048576 --skew 4 --rebalance > gpurun_out/bench_n${N}_skew_rebalance.json 2> gpurun_out/bench_n${N}_skew_rebalance.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus $N --steps 1 --warmup 0 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err
tail -c 600 gpurun_out/bench_n$N.json; tail -c 400 gpurun_out/bench_n${N}_skew.json; tail -c 400 gpurun_out/bench_n${N}_skew_rebalance.json; tail -5 gpurun_out/bench_n${N}_skew_rebalance.err
