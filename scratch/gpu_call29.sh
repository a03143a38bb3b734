#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c29; mkdir -p $O
run() { name=$1; np=$2; shift 2
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@" > $O/$name.out 2> $O/$name.err; echo "$name rc=$?" >> $O/rc.txt; }
run bench8 8 bench.py --gpus 8 --steps 300 --warmup 20
run bench4 4 bench.py --gpus 4 --steps 300 --warmup 20
run bench2 2 bench.py --gpus 2 --steps 300 --warmup 20
run bench8_b1024 8 bench.py --gpus 8 --steps 100 --warmup 10 --batch 1024
run check8 8 tests/multi_gpu_check.py
cat $O/rc.txt
for f in bench2 bench4 bench8 bench8_b1024; do tail -1 $O/$f.out | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), round(d.get('value_hot_l2',0)), 'e2e', round(d['e2e']['value']), d['config'].get('exchange'), d['clocks'].get('reasons'))
except Exception as e: print('$f ERR', e)"; done
grep -v "^\*\|OMP\|^$\|NCCL" $O/check8.out | tail -4
