# two-GPU run (gpurun --gpus 2): the C app's -g 2 path, both NCCL frame gathers against the single-GPU frame,
# bench at N = 2 with the strong legs, raster bench at N = 2
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
nvidia-smi -L > $O/n2_gpus.txt
(timeout 1200 python -m pytest tests/test_gpu_multi.py -m gpu -q -x -rs 2>&1 | tail -12) > $O/n2_pytest.log 2>&1; cat $O/n2_pytest.log
(cd /tmp && timeout 300 $GRAFT_REPO_ROOT/skybox_rt_b200/bin/raytrace -s terrain -n 250 -q 200 -w 1024 -h 1024 -m 2 -o /tmp/one.png && \
 timeout 300 $GRAFT_REPO_ROOT/skybox_rt_b200/bin/raytrace -s terrain -n 250 -q 200 -w 1024 -h 1024 -m 2 -g 2 -o /tmp/two.png -r /tmp/one.png) > $O/n2_capp.log 2>&1; tail -4 $O/n2_capp.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 3 > $O/n2_bench.log 2>&1
tail -1 $O/n2_bench.log | cut -c1-600
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29612 bench.py --gpus 2 --workload draw3d > $O/n2_bench_draw3d.log 2>&1
tail -1 $O/n2_bench_draw3d.log | cut -c1-600
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29613 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 > $O/n2_bench_ref.log 2>&1
tail -1 $O/n2_bench_ref.log | cut -c1-300
