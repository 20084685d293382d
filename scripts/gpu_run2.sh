# GPU-box routine (second half of round 1): smoke, parity tests, full bench, launch list, full ncu captures of the
# solve, thermo and transport kernels (each only after the same command exited 0 without ncu)
mkdir -p gpurun_out
python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python bench.py ${BENCH_ARGS:-} > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err
B="python bench.py --cells 131072 --steps 1 --warmup 1 --no-cpu-baseline"
if [ -z "$SKIP_NCU" ]; then
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches.csv $B > gpurun_out/ncu1.log 2>&1
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:solve_kernel -c 1 -o gpurun_out/prof_solve $B > gpurun_out/ncu2.log 2>&1
BT="python bench.py --cells 4194304 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only"
$BT > gpurun_out/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:thermo_kernel -s 6 -c 2 -o gpurun_out/prof_thermo $BT > gpurun_out/ncu3.log 2>&1
BT2="python bench.py --cells 1048576 --steps 1 --warmup 1 --no-cpu-baseline --thermo-only"
$BT2 > gpurun_out/plain4.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:transport_kernel -s 1 -c 1 -o gpurun_out/prof_transport $BT2 > gpurun_out/ncu4.log 2>&1
fi
tail -3 gpurun_out/smoke.log; tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_full.json
