# round-2 run 9: cooperative repack (tests + timing), clock sampler check, GPU test-suite
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -8) > $O/r9_pytest.log 2>&1; tail -3 $O/r9_pytest.log
timeout 900 python scripts/repack_time.py > $O/r9_repack_time.txt 2>&1; cat $O/r9_repack_time.txt
timeout 900 python bench.py > $O/r9_bench_c3.log 2>&1; tail -1 $O/r9_bench_c3.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['clocks']); print({k:v['clocks'] for k,v in d['strong'].items()})"
timeout 900 python bench.py --workload draw3d > $O/r9_bench_draw3d.log 2>&1; tail -1 $O/r9_bench_draw3d.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print(d['value'], d['ms_per_step'], d['clocks'])"
