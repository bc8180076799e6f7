cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r1_gpu.txt
(timeout 1500 python -m pytest tests -m gpu -q -x 2>&1 | tail -15) > gpurun_out/r1_pytest.log 2>&1
F="python scripts/frame_once.py --workload c3 --frames 4"
( for v in 0 1 2 4 5 6; do timeout 300 $F --variant $v 2>&1 | grep -A1 variant; done
  for v in 0 4 6; do timeout 300 $F --variant $v --stats 1 2>&1 | grep -A1 variant; done
  for v in 0 1 3 4 6; do timeout 300 python scripts/frame_once.py --workload c5 --variant $v --frames 2 2>&1 | grep variant; done
  for v in 0 1 4 6; do timeout 300 python scripts/frame_once.py --workload c2 --variant $v --frames 3 2>&1 | grep variant; done
) > gpurun_out/r1_variants.log 2>&1
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/r1_bench.log 2>&1
tail -3 gpurun_out/r1_pytest.log; cat gpurun_out/r1_variants.log; tail -2 gpurun_out/r1_bench.log
