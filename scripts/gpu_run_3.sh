# round-2 A/B sweep: experiment builds of the trace kernels (python -m skybox_rt_b200.build --exp TAG -D...), refill
# threshold sweep, then the GPU test-suite and the default bench of the tree as it is
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
# A/B of the trace-kernel experiments (lib/exp_*): C3 frame (variant 4), C5 (variants 7, 4), C2
( for e in r2a nosort spec st8 mb5 mb3 mb6st8 specst8; do
    L=skybox_rt_b200/lib/exp_$e/libvortex.so; [ -f $L ] || continue
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c3 --variant 4 --frames 5 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c2 --variant 4 --frames 5 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c5 --variant 7 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c5 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
  done
  for t in 12 16 24 28; do timeout 300 $FO --workload c3 --variant 4 --frames 5 --tune $t 2>&1 | grep variant | sed "s/^/cur tune$t /"; done
  timeout 300 $FO --workload c3 --variant 4 --frames 5 2>&1 | grep variant | sed "s/^/cur /"
  timeout 300 $FO --workload c3 --variant 4 --frames 3 --stats 1 2>&1 | grep -A1 variant | sed "s/^/cur /"
  timeout 300 $FO --workload c2 --variant 4 --frames 5 2>&1 | grep variant | sed "s/^/cur /"
  timeout 300 $FO --workload c5 --variant 7 --frames 3 2>&1 | grep variant | sed "s/^/cur /"
  timeout 300 $FO --workload c5 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/cur /"
) > $O/r3_experiments.log 2>&1
cat $O/r3_experiments.log
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > $O/r3_pytest.log 2>&1; tail -3 $O/r3_pytest.log
timeout 900 python bench.py > $O/r3_bench_c3.log 2>&1; tail -1 $O/r3_bench_c3.log | cut -c1-300
