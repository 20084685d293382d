# the other BASELINE.json configs as bench lines (parity for them is in tests/): H2 1M cells, C7H16 one-step 1M cells,
# drm19 (the C5 flame mechanism) 1M cells; then an N-GPU weak-scaling line if N > 1
N=${1:-1}
mkdir -p gpurun_out
python bench.py --mech es80 --cells 1048576 --steps 3 --warmup 3 > gpurun_out/bench_es80_1M.json 2> gpurun_out/bench_es80.err
python bench.py --mech c7_onestep --cells 1048576 --steps 3 --warmup 3 > gpurun_out/bench_c7_1M.json 2> gpurun_out/bench_c7.err
python bench.py --mech drm19 --cells 1048576 --steps 3 --warmup 3 > gpurun_out/bench_drm19_1M.json 2> gpurun_out/bench_drm19.err
if [ "$N" -gt 1 ]; then
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 2 --warmup 3 --no-thermo > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err
fi
for f in es80_1M c7_1M drm19_1M n$N; do [ -f gpurun_out/bench_$f.json ] && head -c 300 gpurun_out/bench_$f.json && echo; done
