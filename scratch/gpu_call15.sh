#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c15; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu -x -s > $O/multi.log 2>&1; echo "multi rc=$?" >> $O/rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 200 --warmup 10 > $O/bench_2gpu.json 2> $O/bench_2gpu.err; echo "bench2 rc=$?" >> $O/rc.txt
cat $O/rc.txt; grep -v "^\*\|OMP\|^$" $O/multi.log | tail -14; python -c "
import json; d=json.load(open('$O/bench_2gpu.json')); print(d['value'], d['ms_per_step'], d['value_hot_l2'], d['e2e']['value'], d['config']['exchange'])"
