# round 2, GPU run 9: the allocation-free entry point (alone in a fresh process, then at the end of the full GPU
# suite), smoke, and the default bench line
mkdir -p gpurun_out/r2_run9
O=gpurun_out/r2_run9
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/pytest_no_alloc_alone.log 2>&1; echo "pytest rc=$?" >> $O/pytest_no_alloc_alone.log
tail -5 $O/pytest_no_alloc_alone.log
python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
tail -8 $O/pytest_gpu.log
python __graft_entry__.py --smoke > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log; tail -2 $O/smoke.log
( time python bench.py ) > $O/bench_default.json 2> $O/bench_default.err; tail -3 $O/bench_default.err
tail -c 1500 $O/bench_default.json
