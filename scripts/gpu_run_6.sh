# round-2 run 6: raster path (tests + bench with the pipelined e2e loop), pushed-child prefetch A/B on the HBM-resident
# configs, C2 after the chain threshold change
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
(timeout 1200 python -m pytest tests/test_gpu_draw3d.py tests/test_gpu_rt.py -m gpu -q -x 2>&1 | tail -8) > $O/r6_pytest.log 2>&1; tail -3 $O/r6_pytest.log
( for e in cur pf; do
    if [ $e = cur ]; then L=skybox_rt_b200/lib/libvortex.so; else L=skybox_rt_b200/lib/exp_$e/libvortex.so; fi
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c5 --variant 0 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c5 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c4 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c3 --variant 4 --frames 5 2>&1 | grep variant | sed "s/^/$e /"
  done
  timeout 300 $FO --workload c2 --variant 4 --frames 6 2>&1 | grep variant
  timeout 300 $FO --workload c2 --variant 4 --frames 6 --chains 2 2>&1 | grep variant | sed "s/^/chains2 /"
) > $O/r6_experiments.log 2>&1; cat $O/r6_experiments.log
timeout 900 python bench.py --workload draw3d > $O/r6_bench_draw3d.log 2>&1; tail -1 $O/r6_bench_draw3d.log | cut -c1-1500
timeout 900 python bench.py --workload draw3d --size 1024 > $O/r6_bench_draw3d_1024.log 2>&1; tail -1 $O/r6_bench_draw3d_1024.log | cut -c1-600
