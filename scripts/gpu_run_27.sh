# two-GPU box: the multi-GPU tests and the raster / RT suites that upload (shadow path restored), RT bench e2e
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_rt.py tests/test_zz_no_alloc.py -m gpu -q -x -rs 2>&1 | tail -6) > $O/r27_pytest.log 2>&1; cat $O/r27_pytest.log
timeout 400 python bench.py --no-strong > $O/r27_bench_c3.json 2> $O/r27_bench_c3.err; python - <<PY
import json
d=json.loads(open("$O/r27_bench_c3.json").read().strip().splitlines()[-1])
print("C3 value", round(d["value"],1), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), "e2e ms", d["e2e"].get("ms_per_step"), "parity", d["parity_check"]["mismatches"])
PY
