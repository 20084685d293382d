# round 2, GPU run 17: full ncu capture of the final lane-per-cell kernel on es80 (1,048,576 cells) with source import
mkdir -p gpurun_out/r2_run17
O=gpurun_out/r2_run17
G="python bench.py --mech es80 --cells 1048576 --steps 1 --warmup 1 --no-thermo --no-cpu-baseline --no-parity --no-strong --no-flow"
$G > $O/plain.json 2> $O/plain.err && ncu --set full --clock-control none --import-source on -k regex:lpc_solve_kernel -c 1 -o $O/prof_lpc_es80 $G > $O/ncu.log 2>&1
tail -3 $O/ncu.log; tail -c 600 $O/plain.json
