# round-2 run 4: 256-bit node loads A/B, chain-count A/B, ncu of the wavefront kernels with the per-phase budget,
# GPU test-suite, default bench
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
( for e in cur ldg128 l256mb3; do
    if [ $e = cur ]; then L=skybox_rt_b200/lib/libvortex.so; else L=skybox_rt_b200/lib/exp_$e/libvortex.so; fi
    [ -f $L ] || continue
    for c in 1 2; do SKYBOX_RT_LIB=$L timeout 300 $FO --workload c3 --variant 4 --frames 6 --chains $c 2>&1 | grep variant | sed "s/^/$e chains$c /"; done
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c2 --variant 4 --frames 6 --chains 1 2>&1 | grep variant | sed "s/^/$e chains1 /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c5 --variant 7 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c5 --variant 4 --frames 3 2>&1 | grep variant | sed "s/^/$e /"
    SKYBOX_RT_LIB=$L timeout 300 $FO --workload c4 --variant 4 --frames 2 --chains 1 2>&1 | grep variant | sed "s/^/$e chains1 /"
  done
  for c in 3 4; do timeout 300 $FO --workload c3 --variant 4 --frames 6 --chains $c 2>&1 | grep variant | sed "s/^/cur chains$c /"; done
  timeout 300 $FO --workload c2 --variant 4 --frames 6 --chains 2 2>&1 | grep variant | sed "s/^/cur chains2 /"
  timeout 300 $FO --workload c4 --variant 4 --frames 2 --chains 2 2>&1 | grep variant | sed "s/^/cur chains2 /"
  for v in 7 1 2 5; do timeout 300 $FO --workload c3 --variant $v --frames 4 2>&1 | grep variant | sed "s/^/cur /"; done
) > $O/r4_experiments.log 2>&1
cat $O/r4_experiments.log
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > $O/r4_pytest.log 2>&1; tail -3 $O/r4_pytest.log
NCU="ncu --set full --clock-control none --import-source on"
timeout 900 $FO --workload c3 --variant 4 --frames 4 --chains 1 > $O/r4_c3_plain.log 2>&1 && \
timeout 1500 $NCU -k regex:rt_wave -s 12 -c 6 -f -o $O/r4_c3_wave $FO --workload c3 --variant 4 --frames 4 --chains 1 > $O/r4_c3_wave_ncu.log 2>&1
python profiles/ncu_summary.py $O/r4_c3_wave.ncu-rep > $O/r4_c3_wave_summary.txt 2>/dev/null
python profiles/ncu_lines.py $O/r4_c3_wave.ncu-rep 50 -k "rt_wave_trace_kernel<\(int\)0" --budget 2073600 > $O/r4_c3_wave_k1_lines.txt 2>/dev/null
python profiles/ncu_lines.py $O/r4_c3_wave.ncu-rep 50 -k "rt_wave_trace_kernel<\(int\)1" --budget 3194941 > $O/r4_c3_wave_k3_lines.txt 2>/dev/null
python profiles/ncu_lines.py $O/r4_c3_wave.ncu-rep 50 -k "rt_wave_trace_kernel<\(int\)2" > $O/r4_c3_wave_k5_lines.txt 2>/dev/null
ncu -i $O/r4_c3_wave.ncu-rep --page raw --csv > $O/r4_c3_wave_raw.csv 2>/dev/null
rm -f $O/r4_c3_wave.ncu-rep
timeout 900 python bench.py > $O/r4_bench_c3.log 2>&1; tail -1 $O/r4_bench_c3.log | cut -c1-300
