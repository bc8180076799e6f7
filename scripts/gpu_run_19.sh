# byte-granular staging ring: upload stress test, raster + multi tests, e2e timeline and the draw3d bench
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_draw3d.py tests/test_gpu_rt.py -m gpu -q -x 2>&1 | tail -8) > $O/r19_pytest.log 2>&1; cat $O/r19_pytest.log
python scripts/draw3d_e2e_timeline.py > $O/r19_timeline.txt 2>&1; head -2 $O/r19_timeline.txt; tail -5 $O/r19_timeline.txt; VX_CUDA_STAGE_KERNEL=0 python scripts/draw3d_e2e_timeline.py > $O/r19_timeline_dma.txt 2>&1; tail -3 $O/r19_timeline_dma.txt
timeout 300 python bench.py --workload draw3d --steps 30 --warmup 5 > $O/r19_bench_draw3d.json 2> $O/r19_bench_draw3d.err; python - <<PY
import json
d=json.loads(open("$O/r19_bench_draw3d.json").read().strip().splitlines()[-1])
e=d["e2e"]; print("kernel ms", round(d["ms_per_step"],4), "e2e ms", round(e["ms_per_step"],4), "serial", round(e["serial"]["ms_per_step"],3), e["serial"].get("host_ms"))
PY
