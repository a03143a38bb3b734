#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c27; mkdir -p $O
run() { name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus 2 --steps 300 --warmup 20 > $O/$name.out 2> $O/$name.err; echo "$name rc=$?" >> $O/rc.txt
  tail -1 $O/$name.out | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$name', round(d['value']), round(d['ms_per_step'],4), round(d['value_hot_l2']), 'e2e', round(d['e2e']['value']))"; }
run base
B200_WGRAD2_SPLITS=14 B200_WGRAD1_CTAS=64 run co_14_64
B200_WGRAD2_SPLITS=12 B200_WGRAD1_CTAS=64 run co_12_64
B200_WGRAD2_SPLITS=14 B200_WGRAD1_CTAS=56 run co_14_56
cat $O/rc.txt
