# round-2 run 5: fused-sample wavefront passes (C4), per-kernel load width; GPU test-suite; default bench with strong legs
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > $O/r5_pytest.log 2>&1; tail -3 $O/r5_pytest.log
( for c in 0 1 2; do timeout 300 $FO --workload c4 --variant 4 --frames 3 --chains $c 2>&1 | grep variant | sed "s/^/chains$c /"; done
  timeout 300 $FO --workload c4 --variant 7 --frames 2 2>&1 | grep variant
  for c in 1 2; do timeout 300 $FO --workload c3 --variant 4 --frames 6 --chains $c 2>&1 | grep variant | sed "s/^/chains$c /"; done
  timeout 300 $FO --workload c2 --variant 4 --frames 6 2>&1 | grep variant
  timeout 300 $FO --workload c5 --variant 0 --frames 3 2>&1 | grep variant
  timeout 300 $FO --workload c3 --variant 4 --frames 4 --ranks 8 2>&1 | grep variant | sed "s/^/rank0of8 /"
  timeout 300 $FO --workload c4 --variant 4 --frames 4 --ranks 8 2>&1 | grep variant | sed "s/^/rank0of8 /"
  timeout 300 $FO --workload c4 --variant 7 --frames 4 --ranks 8 2>&1 | grep variant | sed "s/^/rank0of8 /"
) > $O/r5_experiments.log 2>&1; cat $O/r5_experiments.log
timeout 900 python bench.py > $O/r5_bench_c3.log 2>&1; tail -1 $O/r5_bench_c3.log | cut -c1-400
