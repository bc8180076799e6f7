# two-GPU A/B of the weak C3 step: fused / separate bounce launches, one / two chains
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
P=29800
for cfg in "0 0" "1 0" "0 1" "1 1"; do
  set -- $cfg; P=$((P+1))
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $P bench.py --gpus 2 --steps 20 --warmup 3 --no-strong --unfused $1 --chains $2 > $O/n2ab_$1_$2.log 2>&1
  tail -1 $O/n2ab_$1_$2.log | python -c "
import json,sys
d=json.loads(sys.stdin.readline()); print('unfused $1 chains $2', round(d['ms_per_step'],4), 'kernel', d['roofline']['kernel_ms'], 'e2e', round(d['e2e']['ms_per_step'],4))"
done
