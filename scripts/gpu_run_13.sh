# round-2 run 13: SAH rotation passes over the device-built radix tree: trace speed against the host SAH tree, build time
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
(timeout 900 python -m pytest tests/test_gpu_lbvh.py tests/test_gpu_rt_configs.py -m gpu -q -x 2>&1 | tail -5) > $O/r13_pytest.log 2>&1; tail -3 $O/r13_pytest.log
( timeout 300 $FO --workload c3 --variant 4 --frames 5 2>&1 | grep variant | sed "s/^/host-sah /"
  timeout 300 $FO --workload c3 --variant 4 --frames 3 --stats 1 2>&1 | grep -A1 variant | sed "s/^/host-sah /"
  for r in 1 2 3 4 6 9; do
    timeout 300 $FO --workload c3 --variant 4 --frames 5 --builder device --rotate $r 2>&1 | grep variant | sed "s/^/lbvh rotate$r /"
    timeout 300 $FO --workload c3 --variant 4 --frames 3 --stats 1 --builder device --rotate $r 2>&1 | grep -A1 variant | sed "s/^/lbvh rotate$r /"
  done
  timeout 300 $FO --workload c2 --variant 4 --frames 5 2>&1 | grep variant | sed "s/^/host-sah /"
  for r in 1 3; do timeout 300 $FO --workload c2 --variant 4 --frames 5 --builder device --rotate $r 2>&1 | grep variant | sed "s/^/lbvh rotate$r /"; done
  timeout 300 $FO --workload c5 --variant 0 --frames 3 2>&1 | grep variant | sed "s/^/host-sah /"
  for r in 1 3; do timeout 300 $FO --workload c5 --variant 0 --frames 3 --builder device --rotate $r 2>&1 | grep variant | sed "s/^/lbvh rotate$r /"; done
) > $O/r13_experiments.log 2>&1; cat $O/r13_experiments.log
timeout 600 python profiles/lbvh_build_time.py > $O/r13_lbvh_build_time.txt 2>&1; cat $O/r13_lbvh_build_time.txt
