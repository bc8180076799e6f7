# host binning pool under SRT_DRAW3D_AUTO: raster tests + the draw3d bench (e2e before: 1.98-2.27 ms per 4096^2 frame)
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 900 python -m pytest tests/test_gpu_draw3d.py tests/test_gpu_units.py -m gpu -q -x 2>&1 | tail -15) > $O/r17_pytest.log 2>&1; cat $O/r17_pytest.log
for i in 1 2; do timeout 300 python bench.py --workload draw3d --steps 30 --warmup 5 > $O/r17_bench_draw3d_$i.json 2> $O/r17_bench_draw3d_$i.err; python - <<PY
import json
d=json.loads(open("$O/r17_bench_draw3d_$i.json").read().strip().splitlines()[-1])
e=d["e2e"]; print("run $i kernel ms", round(d["ms_per_step"],4), "e2e ms", round(e["ms_per_step"],4), "serial", round(e["serial"]["ms_per_step"],3), e["serial"].get("host_ms"))
PY
done
