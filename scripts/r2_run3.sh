# round 2, GPU run 3 (2 GPUs): NCCL rebalancer tests, N=2 bench with the strong-scaled leg, then the 1-GPU default line
mkdir -p gpurun_out/r2_run3
O=gpurun_out/r2_run3
python -m pytest tests/test_gpu_balance.py -m gpu -q -s > $O/pytest_balance.log 2>&1; echo "pytest rc=$?" >> $O/pytest_balance.log
tail -6 $O/pytest_balance.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 2 --warmup 3 --no-thermo > $O/bench_n2.json 2> $O/bench_n2.err; echo "bench n2 rc=$?"
tail -c 2500 $O/bench_n2.json; tail -5 $O/bench_n2.err
python bench.py --steps 2 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench n1 rc=$?"
tail -c 6000 $O/bench_n1.json; tail -5 $O/bench_n1.err
