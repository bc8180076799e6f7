# eight-GPU run (gpurun --gpus 8): the driver's scaling command at N = 8 (weak C3 + strong C4 / C5 legs) and the
# reference arm under torchrun; optional second argument "raster" adds the raster bench
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
N=${1:-8}
nvidia-smi -L > $O/n${N}_gpus.txt
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus $N --steps 20 --warmup 3 > $O/n${N}_bench.log 2>&1
tail -1 $O/n${N}_bench.log | cut -c1-400
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29713 bench.py --gpus $N --impl reference --steps 2 --warmup 1 > $O/n${N}_bench_ref.log 2>&1
tail -1 $O/n${N}_bench_ref.log | cut -c1-200
if [ "$2" = "raster" ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus $N --workload draw3d > $O/n${N}_bench_draw3d.log 2>&1
  tail -1 $O/n${N}_bench_draw3d.log | cut -c1-300
fi
