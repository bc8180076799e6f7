# last validation of the tree as committed: GPU test-suite, smoke, default bench (with its wall time)
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > $O/r15_pytest.log 2>&1; tail -3 $O/r15_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r15_smoke.log 2>&1; tail -2 $O/r15_smoke.log
t0=$(date +%s.%N); python bench.py > $O/r15_bench_c3.log 2>&1; t1=$(date +%s.%N); echo "bench wall $(echo "$t1 - $t0" | bc) s" | tee $O/r15_bench_wall.txt
tail -1 $O/r15_bench_c3.log | cut -c1-300
python bench.py --workload c2 --no-strong > $O/r15_bench_c2.log 2>&1; tail -1 $O/r15_bench_c2.log | cut -c1-200
