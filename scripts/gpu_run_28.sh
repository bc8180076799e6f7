# two ranks under torchrun, as the driver launches the scaling run: RT bench (weak + strong legs) and the raster bench
cd $GRAFT_REPO_ROOT
O=gpurun_out; mkdir -p $O
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29731 bench.py --gpus 2 --steps 10 --warmup 3 > $O/r28_bench_n2.json 2> $O/r28_bench_n2.err; echo "rc=$?"; python - <<PY
import json
d=json.loads(open("$O/r28_bench_n2.json").read().strip().splitlines()[-1])
print("N=2 value", round(d["value"],1), "ms", round(d["ms_per_step"],4), "e2e", round(d["e2e"]["value"],1), "parity", d.get("parity_check",{}).get("mismatches"), "strong", {k:(round(v.get("ms_per_step",0),3), round(v.get("speedup_vs_n1",0),2)) for k,v in d.get("strong",{}).items() if isinstance(v,dict)}, "clocks", d["clocks"].get("samples"))
PY
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29732 bench.py --gpus 2 --workload draw3d --steps 10 --warmup 3 > $O/r28_bench_draw3d_n2.json 2> $O/r28_bench_draw3d_n2.err; echo "rc=$?"; tail -c 600 $O/r28_bench_draw3d_n2.json
