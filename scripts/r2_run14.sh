# round 2, GPU run 14: full ncu capture of thermo_kernel (gri30, 4,194,304 cells) with source import, for the per-line table
mkdir -p gpurun_out/r2_run14
O=gpurun_out/r2_run14
G="python bench.py --mech gri30 --thermo-only --no-cpu-baseline"
$G > $O/plain.json 2> $O/plain.err && ncu --set full --clock-control none --import-source on -k regex:thermo_kernel -c 2 -o $O/prof_thermo $G > $O/ncu.log 2>&1
tail -3 $O/ncu.log; tail -c 1200 $O/plain.json
