# round-2 FINAL evidence run (kernels as shipped): GPU test-suite, ncu --set full of the shipped kernels on C2 / C3 /
# C4 / C5 with per-phase instruction budgets (computed here so that line numbers match the build), launch list of
# the default C3 frame, bench lines of every config (never under ncu), raster bench, reference arm
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
NCU="ncu --set full --clock-control none --import-source on"
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > $O/r12_pytest.log 2>&1; tail -3 $O/r12_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r12_smoke.log 2>&1; tail -2 $O/r12_smoke.log
cap() { # name, ncu filter args..., -- command
  name=$1; shift
  filt=(); while [ "$1" != "--" ]; do filt+=("$1"); shift; done; shift
  timeout 900 "$@" > $O/${name}_plain.log 2>&1 || { echo "$name: plain run failed"; return; }
  timeout 1500 $NCU "${filt[@]}" -f -o $O/$name "$@" > $O/${name}_ncu.log 2>&1
  python profiles/ncu_summary.py $O/$name.ncu-rep > $O/${name}_summary.txt 2>/dev/null
}
$FO --workload c3 --variant 0 --frames 4 > $O/r02c_c3_plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02c_c3_launches.csv $FO --workload c3 --variant 0 --frames 4 > /dev/null 2>&1
# one chain for the capture: the kernels of a frame in program order (two chains interleave two streams)
cap r02c_c3_wave -k regex:rt_wave -s 8 -c 4 -- $FO --workload c3 --variant 4 --frames 4 --chains 1
python profiles/ncu_lines.py $O/r02c_c3_wave.ncu-rep 45 -k rt_wave_trace_kernel --skip 0 --budget 2073600 > $O/r02c_c3_wave_k1_lines.txt 2>/dev/null
python profiles/ncu_lines.py $O/r02c_c3_wave.ncu-rep 45 -k rt_wave_trace_kernel --skip 1 --budget 3194941 > $O/r02c_c3_wave_k3_lines.txt 2>/dev/null
python profiles/ncu_lines.py $O/r02c_c3_wave.ncu-rep 30 -k rt_wave_shade1_kernel --budget 2073600 > $O/r02c_c3_wave_shade1_lines.txt 2>/dev/null
ncu -i $O/r02c_c3_wave.ncu-rep --page raw --csv > $O/r02c_c3_wave_raw.csv 2>/dev/null
cap r02c_c2_wave -k regex:rt_wave -s 6 -c 3 -- $FO --workload c2 --variant 4 --frames 4
python profiles/ncu_lines.py $O/r02c_c2_wave.ncu-rep 45 -k rt_wave_trace_kernel --budget 1048576 > $O/r02c_c2_wave_k1_lines.txt 2>/dev/null
rm -f $O/r02c_c2_wave.ncu-rep
cap r02c_c4_wave -k regex:rt_wave_trace -s 2 -c 2 -- $FO --workload c4 --variant 4 --frames 2 --chains 1
rm -f $O/r02c_c4_wave.ncu-rep
cap r02c_c5_vote -k regex:rays_kernel -s 1 -c 1 -- $FO --workload c5 --variant 0 --frames 2
python profiles/ncu_lines.py $O/r02c_c5_vote.ncu-rep 45 --budget 16777216 > $O/r02c_c5_vote_lines.txt 2>/dev/null
rm -f $O/r02c_c5_vote.ncu-rep
timeout 900 python bench.py > $O/r12_bench_c3.log 2>&1
timeout 900 python bench.py --workload c2 > $O/r12_bench_c2.log 2>&1
timeout 900 python bench.py --workload c1 > $O/r12_bench_c1.log 2>&1
timeout 1200 python bench.py --workload c5 --steps 5 > $O/r12_bench_c5.log 2>&1
timeout 1200 python bench.py --workload c4 --steps 5 > $O/r12_bench_c4.log 2>&1
timeout 900 python bench.py --workload draw3d > $O/r12_bench_draw3d.log 2>&1
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/r12_bench_ref.log 2>&1
timeout 900 python bench.py --builder device --no-strong --no-cpu-baseline > $O/r12_bench_c3_devbuild.log 2>&1
for f in c3 c2 c1 c5 c4 draw3d ref c3_devbuild; do tail -1 $O/r12_bench_$f.log | cut -c1-200; done
du -sh $O
