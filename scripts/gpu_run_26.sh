# validation of the committed tree on one GPU: whole GPU suite, smoke, default bench (C3) and the raster bench
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 900 python -m pytest tests -m gpu -q -x -rs 2>&1 | tail -8) > $O/r26_pytest.log 2>&1; cat $O/r26_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/r26_smoke.log 2>&1; tail -2 $O/r26_smoke.log
timeout 600 python bench.py > $O/r26_bench_c3.json 2> $O/r26_bench_c3.err; python - <<PY
import json
d=json.loads(open("$O/r26_bench_c3.json").read().strip().splitlines()[-1])
print("C3 value", round(d["value"],1), d["unit"], "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), "e2e ms", d["e2e"].get("ms_per_step"), "parity", d.get("parity_check"), "launches", d.get("gpu_launches"))
print("roofline frac", d["roofline"].get("frac"), "cpu", d["cpu_baseline"].get("value"), "variants", d.get("variants_ms"))
PY
timeout 300 python bench.py --workload draw3d > $O/r26_bench_draw3d.json 2> $O/r26_bench_draw3d.err; python - <<PY
import json
d=json.loads(open("$O/r26_bench_draw3d.json").read().strip().splitlines()[-1])
e=d["e2e"]; print("draw3d kernel ms", round(d["ms_per_step"],4), "e2e ms", round(e["ms_per_step"],4), "serial", round(e["serial"]["ms_per_step"],3))
PY
