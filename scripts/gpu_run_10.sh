# round-2 run 10: bounce shading + second shadow rays folded into the shadow/bounce trace launch (A/B against the
# separate K4 / K5 launches), GPU test-suite
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > $O/r10_pytest.log 2>&1; tail -3 $O/r10_pytest.log
( for u in 0 1; do
    for c in 1 2; do timeout 300 $FO --workload c3 --variant 4 --frames 6 --chains $c --unfused $u 2>&1 | grep variant | sed "s/^/unfused$u chains$c /"; done
    timeout 300 $FO --workload c3 --variant 4 --frames 3 --stats 1 --unfused $u 2>&1 | grep -A1 variant | sed "s/^/unfused$u /"
    timeout 300 $FO --workload c4 --variant 4 --frames 3 --unfused $u 2>&1 | grep variant | sed "s/^/unfused$u /"
    timeout 300 $FO --workload c4 --variant 4 --frames 4 --ranks 8 --unfused $u 2>&1 | grep variant | sed "s/^/unfused$u rank0of8 /"
    timeout 300 $FO --workload c3 --variant 5 --frames 4 --unfused $u 2>&1 | grep variant | sed "s/^/unfused$u /"
  done
) > $O/r10_experiments.log 2>&1; cat $O/r10_experiments.log
timeout 900 python bench.py > $O/r10_bench_c3.log 2>&1; tail -1 $O/r10_bench_c3.log | cut -c1-300
