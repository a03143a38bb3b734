#!/bin/bash
# final multi-GPU measurements (round 2): 2 / 4 / 8 workers, batch 128 and the large batches, reference arm, equivalence check
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c23; mkdir -p $O
run() { name=$1; np=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@" > $O/$name.out 2> $O/$name.err; echo "$name rc=$?" >> $O/rc.txt; }
run check8 8 tests/multi_gpu_check.py
run check2 2 tests/multi_gpu_check.py
run bench2 2 bench.py --gpus 2 --steps 300 --warmup 20
run bench4 4 bench.py --gpus 4 --steps 300 --warmup 20
run bench8 8 bench.py --gpus 8 --steps 300 --warmup 20
run bench8_b512 8 bench.py --gpus 8 --steps 100 --warmup 10 --batch 512
run bench8_b1024 8 bench.py --gpus 8 --steps 100 --warmup 10 --batch 1024
run bench8_b4096 8 bench.py --gpus 8 --steps 50 --warmup 5 --batch 4096
run bench8_wide 8 bench.py --gpus 8 --steps 50 --warmup 5 --model wide
run ref8 8 bench.py --impl reference --gpus 8 --steps 64 --warmup 3
run ref2 2 bench.py --impl reference --gpus 2 --steps 100 --warmup 3
timeout 600 python -m pytest tests/test_gpu_multi.py -q -m gpu -s > $O/multi_pytest.log 2>&1; echo "multi pytest rc=$?" >> $O/rc.txt
cat $O/rc.txt
grep -v "^\*\|OMP\|^$\|NCCL" $O/check8.out | tail -5
for f in bench2 bench4 bench8 bench8_b512 bench8_b1024 bench8_b4096 bench8_wide ref8 ref2; do tail -1 $O/$f.out | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), round(d.get('value_hot_l2',0)), round(d['e2e']['value']), d['config'].get('exchange'), d.get('clocks',{}) and d['clocks'].get('reasons'), d['e2e'].get('last_loss'))
except Exception as e: print('$f ERR', e)"; done
tail -3 $O/multi_pytest.log
