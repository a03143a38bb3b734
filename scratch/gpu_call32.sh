#!/bin/bash
# 4-GPU box: one-shot push exchange of the conv gradients at world 4 and 2, against the two-barrier kernel
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c32; mkdir -p $O
export B200_P2P_TIMEOUT_S=8
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 150 $TR --nproc-per-node 4 --master-port 29761 tests/multi_gpu_check.py > $O/check4.log 2>&1; echo "check4 rc=$?" >> $O/rc.txt
grep -v "^\[W\|NCCL\|^W1" $O/check4.log | tail -6
if grep -q "check4 rc=0" $O/rc.txt; then
timeout 120 $TR --nproc-per-node 4 --master-port 29762 bench.py --gpus 4 --steps 300 --warmup 20 > $O/bench4_push.json 2> $O/bench4_push.err; echo "bench4 push rc=$?" >> $O/rc.txt
timeout 120 $TR --nproc-per-node 2 --master-port 29764 tests/multi_gpu_check.py > $O/check2.log 2>&1; echo "check2 rc=$?" >> $O/rc.txt
timeout 120 $TR --nproc-per-node 2 --master-port 29765 bench.py --gpus 2 --steps 300 --warmup 20 > $O/bench2_push.json 2> $O/bench2_push.err; echo "bench2 push rc=$?" >> $O/rc.txt
fi
B200CNN_INBOX=0 timeout 120 $TR --nproc-per-node 4 --master-port 29763 bench.py --gpus 4 --steps 300 --warmup 20 > $O/bench4_old.json 2> $O/bench4_old.err; echo "bench4 old rc=$?" >> $O/rc.txt
cat $O/rc.txt
grep -v "^\[W\|NCCL\|^W1" $O/check2.log | tail -5
for f in bench4_push bench4_old bench2_push; do tail -1 $O/$f.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), d['config'].get('exchange'), 'e2e', round(d['e2e']['value']))" 2>&1 | tail -1; done
tail -3 $O/bench4_push.err
