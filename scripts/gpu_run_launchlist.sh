# launch lists of the committed defaults: scripts/frame_once.py (C3, 4 frames) and bench.py itself (2 timed steps)
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
FO="python scripts/frame_once.py"
$FO --workload c3 --variant 0 --frames 4 > $O/r02d_c3_plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r02d_c3_launches.csv $FO --workload c3 --variant 0 --frames 4 > /dev/null 2>&1
B="python bench.py --steps 2 --warmup 3 --no-strong --no-cpu-baseline --no-variants"
$B > $O/r02d_bench_plain.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/r02d_bench_launches.csv $B > $O/r02d_bench_under_ncu.log 2>&1
tail -1 $O/r02d_bench_plain.log | cut -c1-200; wc -l $O/r02d_*launches.csv
