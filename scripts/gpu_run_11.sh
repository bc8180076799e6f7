# round-2 run 11: speculative traversal with 2 / 3 parked triangles per lane
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
( for e in cur spec2 spec2mb3 spec3mb3; do
    if [ $e = cur ]; then L=skybox_rt_b200/lib/libvortex.so; else L=skybox_rt_b200/lib/exp_$e/libvortex.so; fi
    [ -f $L ] || continue
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c3 --variant 4 --frames 6 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c3 --variant 4 --frames 3 --stats 1 2>&1 | grep -A1 variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c2 --variant 4 --frames 6 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c4 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c5 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
  done
) > $O/r11_experiments.log 2>&1; cat $O/r11_experiments.log
