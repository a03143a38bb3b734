#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c16; mkdir -p $O
nvidia-smi -L | wc -l > $O/ngpu.txt
run() { # name nproc extra-env args...
  name=$1; np=$2; shift 2
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) "$@" > $O/$name.out 2> $O/$name.err; echo "$name rc=$?" >> $O/rc.txt
}
run check8 8 tests/multi_gpu_check.py
run bench8 8 bench.py --gpus 8 --steps 200 --warmup 10
B200CNN_OVERLAP=1 run bench8_ovl 8 bench.py --gpus 8 --steps 200 --warmup 10
run bench4 4 bench.py --gpus 4 --steps 200 --warmup 10
B200CNN_OVERLAP=0 run bench4_single 4 bench.py --gpus 4 --steps 200 --warmup 10
run bench8_b1024 8 bench.py --gpus 8 --steps 100 --warmup 10 --batch 1024
run bench8_b4096 8 bench.py --gpus 8 --steps 50 --warmup 5 --batch 4096
run bench8_b512 8 bench.py --gpus 8 --steps 100 --warmup 10 --batch 512
run ref8 8 bench.py --impl reference --gpus 8 --steps 64 --warmup 3
cat $O/rc.txt
grep -v "^\*\|OMP\|^$\|NCCL" $O/check8.out | tail -6
for f in bench8 bench8_ovl bench4 bench4_single bench8_b512 bench8_b1024 bench8_b4096 ref8; do tail -1 $O/$f.out | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), round(d.get('value_hot_l2',0)), round(d['e2e']['value']), d['config'].get('exchange'), d.get('clocks',{}) and d['clocks'].get('reasons'))
except Exception as e: print('$f ERR', e)"; done
