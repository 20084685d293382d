# round 2, GPU run 30: last check of the rebuilt committed library -- full GPU suite and smoke
mkdir -p gpurun_out/r2_run30
O=gpurun_out/r2_run30
python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -3 $O/pytest_gpu.log
python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log; tail -2 $O/smoke.log
