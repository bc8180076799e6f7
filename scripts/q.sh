M="smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,l1tex__t_sector_hit_rate.pct"
for v in 0 1; do for m in 0 1 2; do
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --variant $v --mode $m > gpurun_out/plain_${v}_${m}.log 2>&1 && ncu --metrics $M --clock-control none -k regex:"rt_(wide|baseline)" -s 3 -c 1 --csv --log-file gpurun_out/q_${v}_${m}.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --variant $v --mode $m > /dev/null 2>&1
done; done
