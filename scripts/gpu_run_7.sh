# round-2 evidence run: GPU test-suite, ncu --set full of the shipped kernels on C2 / C3 / C4 / C5 with per-phase
# instruction budgets (profiles/ncu_lines.py --budget, computed here so that line numbers match the build), launch
# list of the default C3 frame, bench lines of every config (never under ncu), raster bench
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
NCU="ncu --set full --clock-control none --import-source on"
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > $O/r7_pytest.log 2>&1; tail -3 $O/r7_pytest.log
cap() { # name, ncu filter args..., -- command
  name=$1; shift
  filt=(); while [ "$1" != "--" ]; do filt+=("$1"); shift; done; shift
  timeout 900 "$@" > $O/${name}_plain.log 2>&1 || { echo "$name: plain run failed"; return; }
  timeout 1500 $NCU "${filt[@]}" -f -o $O/$name "$@" > $O/${name}_ncu.log 2>&1
  python profiles/ncu_summary.py $O/$name.ncu-rep > $O/${name}_summary.txt 2>/dev/null
}
# launch list of the default C3 frame (2 chains)
$FO --workload c3 --variant 0 --frames 4 > $O/r02b_c3_plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02b_c3_launches.csv $FO --workload c3 --variant 0 --frames 4 > /dev/null 2>&1
cap r02b_c3_wave -k regex:rt_wave -s 12 -c 6 -- $FO --workload c3 --variant 4 --frames 4 --chains 1
python profiles/ncu_lines.py $O/r02b_c3_wave.ncu-rep 45 -k "rt_wave_trace_kernel<\(int\)0" --budget 2073600 > $O/r02b_c3_wave_k1_lines.txt 2>/dev/null
python profiles/ncu_lines.py $O/r02b_c3_wave.ncu-rep 45 -k "rt_wave_trace_kernel<\(int\)1" --budget 3194941 > $O/r02b_c3_wave_k3_lines.txt 2>/dev/null
python profiles/ncu_lines.py $O/r02b_c3_wave.ncu-rep 45 -k "rt_wave_trace_kernel<\(int\)2" --budget 1000000 > $O/r02b_c3_wave_k5_lines.txt 2>/dev/null
rm -f $O/r02b_c3_wave.ncu-rep
cap r02b_c2_wave -k regex:rt_wave -s 6 -c 3 -- $FO --workload c2 --variant 4 --frames 4
python profiles/ncu_lines.py $O/r02b_c2_wave.ncu-rep 45 -k "rt_wave_trace_kernel" --budget 1048576 > $O/r02b_c2_wave_k1_lines.txt 2>/dev/null
rm -f $O/r02b_c2_wave.ncu-rep
cap r02b_c4_wave -k regex:rt_wave_trace -s 3 -c 3 -- $FO --workload c4 --variant 4 --frames 2 --chains 1
rm -f $O/r02b_c4_wave.ncu-rep
cap r02b_c5_vote -k regex:rays_kernel -s 1 -c 1 -- $FO --workload c5 --variant 0 --frames 2
python profiles/ncu_lines.py $O/r02b_c5_vote.ncu-rep 45 --budget 16777216 > $O/r02b_c5_vote_lines.txt 2>/dev/null
rm -f $O/r02b_c5_vote.ncu-rep
# bench lines
timeout 900 python bench.py > $O/r7_bench_c3.log 2>&1
timeout 900 python bench.py --workload c2 > $O/r7_bench_c2.log 2>&1
timeout 1200 python bench.py --workload c5 --steps 5 > $O/r7_bench_c5.log 2>&1
timeout 1200 python bench.py --workload c4 --steps 5 > $O/r7_bench_c4.log 2>&1
timeout 900 python bench.py --workload draw3d > $O/r7_bench_draw3d.log 2>&1
VX_CUDA_STAGED_UPLOADS=0 timeout 900 python bench.py --workload draw3d > $O/r7_bench_draw3d_nostage.log 2>&1
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/r7_bench_ref.log 2>&1
for f in c3 c2 c5 c4 draw3d draw3d_nostage ref; do tail -1 $O/r7_bench_$f.log | cut -c1-200; done
du -sh $O
