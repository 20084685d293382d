# round 2, GPU run 2: the whole GPU suite (new: equal-tolerance, 256-cell goldens, C5) + smoke
mkdir -p gpurun_out/r2_run2
O=gpurun_out/r2_run2
python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
DFCHEM_SOLVER=lane python -m pytest tests/test_gpu_parity.py tests/test_equal_tolerance.py -m gpu -q -k "es80 or h2_li or c7" > $O/pytest_lane.log 2>&1; echo "pytest(lane) rc=$?" >> $O/pytest_lane.log
tail -15 $O/pytest_gpu.log; tail -3 $O/smoke.log; tail -5 $O/pytest_lane.log
