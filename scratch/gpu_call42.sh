#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c42; mkdir -p $O
export B200_P2P_TIMEOUT_S=8
timeout 45 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 8 --master-port 29762 bench.py --gpus 8 --steps 300 --warmup 20 --no-cpu-baseline > $O/bench8.json 2> $O/bench8.err; echo "bench8 rc=$?"
tail -1 $O/bench8.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('bench8', round(d['value']), round(d['ms_per_step'],4), round(d['value_hot_l2']), d['config'].get('exchange'), 'e2e', round(d['e2e']['value']))" 2>&1 | tail -1
