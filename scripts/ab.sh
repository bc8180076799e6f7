#!/bin/bash
# usage: scripts/ab.sh "<exp tags>" "<variants>" [extra bench args]   -- prints Mrays/s per (lib, variant)
tags="$1"; vars="$2"; shift 2
for t in $tags; do
  if [ "$t" = "main" ]; then unset SKYBOX_RT_LIB; else export SKYBOX_RT_LIB=$PWD/skybox_rt_b200/lib/exp_$t/libvortex.so; fi
  for v in $vars; do
    python bench.py --steps 5 --warmup 3 --no-cpu-baseline --variant $v "$@" | python -c "import json,sys; d=json.loads(sys.stdin.read()); print('$t', 'variant', $v, round(d['value'],1), 'Mrays/s', round(d['ms_per_step'],4), 'ms', 'nodes', d['roofline']['nodes_fetched'], 'tris', d['roofline']['tris_tested'])"
  done
done
