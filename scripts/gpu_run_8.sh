# round-2 run 8: speculative traversal without spills (3 CTAs/SM), host-sorted C5 rays, C3 wavefront capture with
# per-kernel per-phase budgets
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
NCU="ncu --set full --clock-control none --import-source on"
( for e in cur mb3 specmb3; do
    if [ $e = cur ]; then L=skybox_rt_b200/lib/libvortex.so; else L=skybox_rt_b200/lib/exp_$e/libvortex.so; fi
    [ -f $L ] || continue
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c3 --variant 4 --frames 6 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c3 --variant 4 --frames 3 --stats 1 2>&1 | grep -A1 variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c2 --variant 4 --frames 6 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c4 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c5 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
  done
  for g in 0 8 16 32; do
    timeout 600 $FO --workload c5 --variant 0 --frames 3 --sort $g 2>&1 | grep variant | sed "s/^/sort$g /"
    timeout 600 $FO --workload c5 --variant 4 --frames 3 --sort $g 2>&1 | grep variant | sed "s/^/sort$g /"
    timeout 600 $FO --workload c5 --variant 1 --frames 3 --sort $g 2>&1 | grep variant | sed "s/^/sort$g /"
  done
  timeout 600 $FO --workload c5 --variant 0 --frames 2 --sort 16 --stats 1 2>&1 | grep -A1 variant | sed "s/^/sort16 /"
  timeout 600 $FO --workload c5 --variant 0 --frames 2 --stats 1 2>&1 | grep -A1 variant | sed "s/^/sort0 /"
) > $O/r8_experiments.log 2>&1; cat $O/r8_experiments.log
timeout 900 $FO --workload c3 --variant 4 --frames 4 --chains 1 > $O/r02b_c3_wave_plain.log 2>&1 && \
timeout 1500 $NCU -k regex:rt_wave -s 12 -c 6 -f -o $O/r02b_c3_wave $FO --workload c3 --variant 4 --frames 4 --chains 1 > $O/r02b_c3_wave_ncu.log 2>&1
python profiles/ncu_summary.py $O/r02b_c3_wave.ncu-rep > $O/r02b_c3_wave_summary.txt 2>/dev/null
python profiles/ncu_lines.py $O/r02b_c3_wave.ncu-rep 45 -k rt_wave_trace_kernel --skip 0 --budget 2073600 > $O/r02b_c3_wave_k1_lines.txt 2>/dev/null
python profiles/ncu_lines.py $O/r02b_c3_wave.ncu-rep 45 -k rt_wave_trace_kernel --skip 1 --budget 3194941 > $O/r02b_c3_wave_k3_lines.txt 2>/dev/null
python profiles/ncu_lines.py $O/r02b_c3_wave.ncu-rep 45 -k rt_wave_trace_kernel --skip 2 --budget 1000000 > $O/r02b_c3_wave_k5_lines.txt 2>/dev/null
python profiles/ncu_lines.py $O/r02b_c3_wave.ncu-rep 30 -k rt_wave_shade1_kernel --budget 2073600 > $O/r02b_c3_wave_shade1_lines.txt 2>/dev/null
head -12 $O/r02b_c3_wave_k3_lines.txt
