#!/bin/bash
# 8-GPU box: push exchange at world 8 (check + bench)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c33; mkdir -p $O
export B200_P2P_TIMEOUT_S=8
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 150 $TR --nproc-per-node 8 --master-port 29761 tests/multi_gpu_check.py > $O/check8.log 2>&1; echo "check8 rc=$?" >> $O/rc.txt
grep -v "^\[W\|NCCL\|^W1" $O/check8.log | tail -6
timeout 120 $TR --nproc-per-node 8 --master-port 29762 bench.py --gpus 8 --steps 300 --warmup 20 > $O/bench8_push.json 2> $O/bench8_push.err; echo "bench8 push rc=$?" >> $O/rc.txt
cat $O/rc.txt
for f in bench8_push; do tail -1 $O/$f.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), d['config'].get('exchange'), 'e2e', round(d['e2e']['value']))" 2>&1 | tail -1; done
tail -3 $O/bench8_push.err
