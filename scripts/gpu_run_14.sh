# round-2 run 14: last validation of the tree as committed (GPU test-suite, smoke, default bench) + the any-hit
# traversal-order experiment
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > $O/r14_pytest.log 2>&1; tail -3 $O/r14_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r14_smoke.log 2>&1; tail -2 $O/r14_smoke.log
( for e in cur anyu; do
    if [ $e = cur ]; then L=skybox_rt_b200/lib/libvortex.so; else L=skybox_rt_b200/lib/exp_$e/libvortex.so; fi
    [ -f $L ] || continue
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c3 --variant 4 --frames 6 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c3 --variant 4 --frames 3 --stats 1 2>&1 | grep -A1 variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c4 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
  done
) > $O/r14_experiments.log 2>&1; cat $O/r14_experiments.log
/usr/bin/time -v python bench.py > $O/r14_bench_c3.log 2> $O/r14_bench_c3.time; tail -1 $O/r14_bench_c3.log | cut -c1-300; grep -E "Elapsed|Maximum resident" $O/r14_bench_c3.time
