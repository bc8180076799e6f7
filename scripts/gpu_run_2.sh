# round-2 ncu evidence: wavefront pipeline on C3 / C2 / C4, ray-list kernels on C5, megakernel + baseline A/B on C3
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
NCU="ncu --set full --clock-control none --import-source on"
cap() { # name, ncu filter args..., -- command
  name=$1; shift
  filt=(); while [ "$1" != "--" ]; do filt+=("$1"); shift; done; shift
  timeout 900 "$@" > $O/${name}_plain.log 2>&1 || { echo "$name: plain run failed"; return; }
  timeout 1500 $NCU "${filt[@]}" -f -o $O/$name "$@" > $O/${name}_ncu.log 2>&1
  python profiles/ncu_summary.py $O/$name.ncu-rep > $O/${name}_summary.txt 2>/dev/null
  python profiles/ncu_lines.py $O/$name.ncu-rep 60 > $O/${name}_lines.txt 2>/dev/null
}
# launch list of the default C3 frame
$FO --workload c3 --variant 0 --frames 4 > $O/r02_c3_plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02_c3_wave_launches.csv $FO --workload c3 --variant 0 --frames 4 > $O/r02_c3_launch_ncu.log 2>&1
cap r02_c3_wave -k regex:rt_wave -s 12 -c 6 -- $FO --workload c3 --variant 4 --frames 4
cap r02_c3_mega_vote -k regex:rt_wide_render -s 2 -c 1 -- $FO --workload c3 --variant 7 --frames 4
cap r02_c3_baseline -k regex:rt_baseline -s 2 -c 1 -- $FO --workload c3 --variant 1 --frames 4
cap r02_c2_wave -k regex:rt_wave -s 6 -c 3 -- $FO --workload c2 --variant 4 --frames 4
cap r02_c5_vote -k regex:rays_kernel -s 1 -c 1 -- $FO --workload c5 --variant 7 --frames 2
cap r02_c5_baseline -k regex:rays_kernel -s 1 -c 1 -- $FO --workload c5 --variant 1 --frames 2
cap r02_c5_pool -k regex:rays_kernel -s 1 -c 1 -- $FO --workload c5 --variant 3 --frames 2
cap r02_c5_while -k regex:rays_kernel -s 1 -c 1 -- $FO --workload c5 --variant 4 --frames 2
cap r02_c4_wave -k regex:rt_wave_trace -s 24 -c 3 -- $FO --workload c4 --variant 4 --frames 2
# bench lines (never under ncu)
timeout 900 python bench.py > $O/r2_bench_c3.log 2>&1
timeout 900 python bench.py --workload c2 > $O/r2_bench_c2.log 2>&1
timeout 1200 python bench.py --workload c5 --steps 5 > $O/r2_bench_c5.log 2>&1
timeout 1200 python bench.py --workload c4 --steps 5 > $O/r2_bench_c4.log 2>&1
ls -la $O/*.ncu-rep; du -sh $O
# keep the transfer under the 64 MiB cap: reports beyond it stay on the box (their summaries travel)
tot=0; for f in $(ls -S -r $O/*.ncu-rep); do s=$(stat -c %s $f); tot=$((tot+s)); if [ $tot -gt 45000000 ]; then echo "dropping $f ($s bytes)"; rm -f $f; fi; done
tail -1 $O/r2_bench_c3.log | cut -c1-600
