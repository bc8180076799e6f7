# packed AUTO frames (one arena upload per frame, draws started back to back): raster tests, e2e timeline, draw3d bench
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 900 python -m pytest tests/test_gpu_draw3d.py tests/test_gpu_units.py -m gpu -q -x 2>&1 | tail -8) > $O/r22_pytest.log 2>&1; cat $O/r22_pytest.log
python scripts/draw3d_e2e_timeline.py > $O/r22_timeline.txt 2>&1; head -4 $O/r22_timeline.txt; tail -6 $O/r22_timeline.txt
timeout 300 python bench.py --workload draw3d --steps 30 --warmup 5 > $O/r22_bench_draw3d.json 2> $O/r22_bench_draw3d.err; python - <<PY
import json
d=json.loads(open("$O/r22_bench_draw3d.json").read().strip().splitlines()[-1])
e=d["e2e"]; print("kernel ms", round(d["ms_per_step"],4), "e2e ms", round(e["ms_per_step"],4), "serial", round(e["serial"]["ms_per_step"],3), e["serial"].get("host_ms"))
PY
