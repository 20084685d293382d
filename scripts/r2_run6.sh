# round 2, GPU run 6: full GPU suite with everything added since run 2 (patches, mock caller, mechanism classes,
# transport fits, C5), then launch list + full ncu capture of the solve kernel and of lpc_solve_kernel
mkdir -p gpurun_out/r2_run6
O=gpurun_out/r2_run6
python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log; tail -2 $O/smoke.log
B="python bench.py --cells 131072 --steps 1 --warmup 1 --no-cpu-baseline --no-configs --no-parity"
$B > $O/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/launches_gri30_131072.csv $B > $O/ncu1.log 2>&1
$B > $O/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:solve_kernel -c 1 -o $O/prof_solve $B > $O/ncu2.log 2>&1
C="python bench.py --mech c7_onestep --cells 1048576 --steps 1 --warmup 1 --no-cpu-baseline --no-thermo"
$C > $O/plain3.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:lpc_solve_kernel -c 1 -o $O/prof_lpc $C > $O/ncu3.log 2>&1
ls -la $O
