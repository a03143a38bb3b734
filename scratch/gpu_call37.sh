#!/bin/bash
# 1 GPU: conv1 filter gradient with 16 worker warps (B200_WGRAD1_WORKERS=256 = the old shape)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c37; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_wide.py -q -m gpu -x -k "conv1 or step or determin or wide" > $O/tests.log 2>&1; echo "tests rc=$?" >> $O/rc.txt
tail -3 $O/tests.log
if grep -q "tests rc=0" $O/rc.txt; then
for nw in 512 256; do
B200_WGRAD1_WORKERS=$nw timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_nw$nw.json 2> $O/bench_nw$nw.err; echo "bench nw$nw rc=$?" >> $O/rc.txt
done
fi
cat $O/rc.txt
for nw in 512 256; do tail -1 $O/bench_nw$nw.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('nw$nw', round(d['value']), round(d['ms_per_step'],4), 'hot', round(d['value_hot_l2']), 'e2e', round(d['e2e']['value']))" 2>&1 | tail -1; done
