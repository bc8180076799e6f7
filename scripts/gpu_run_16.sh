# raster guards (tile size, malformed buffers), unit goldens of tex / raster / om through the raster path, draw3d bench for regressions
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
(timeout 900 python -m pytest tests/test_gpu_units.py tests/test_gpu_draw3d.py tests/test_gpu_binning.py tests/test_gpu_perf.py tests/test_gpu_ref_stub.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -15) > $O/r16_pytest.log 2>&1; cat $O/r16_pytest.log
timeout 300 python bench_draw3d.py --steps 20 --warmup 3 > $O/r16_bench_draw3d.json 2> $O/r16_bench_draw3d.err; tail -c 1500 $O/r16_bench_draw3d.json
