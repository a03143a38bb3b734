#!/bin/bash
# 1 GPU: conv1 forward with 16 worker warps (B200_CONV1_WORKERS=256 = the old shape); schedule trace; wgrad wave split 43 + 21x5
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c38; mkdir -p $O
timeout 300 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_wide.py -q -m gpu -x -k "conv1 or step or determin or wide or 32x32" > $O/tests.log 2>&1; echo "tests rc=$?" >> $O/rc.txt
tail -3 $O/tests.log
if grep -q "tests rc=0" $O/rc.txt; then
B200_CONV1_WORKERS=512 timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_a.json 2> $O/bench_a.err; echo "bench a rc=$?" >> $O/rc.txt
B200_CONV1_WORKERS=256 timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_b.json 2> $O/bench_b.err; echo "bench b rc=$?" >> $O/rc.txt
B200_WGRAD1_CTAS=43 B200_WGRAD2_SPLITS=21 timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_c.json 2> $O/bench_c.err; echo "bench c rc=$?" >> $O/rc.txt
B200CNN_LIB=$PWD/scratch/libb200cnn_trace.so timeout 120 python scratch/trace_step.py 128 > $O/trace.log 2>&1; echo "trace rc=$?" >> $O/rc.txt
fi
cat $O/rc.txt
for f in a b c; do tail -1 $O/bench_$f.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), 'hot', round(d['value_hot_l2']), 'e2e', round(d['e2e']['value']))" 2>&1 | tail -1; done
head -40 $O/trace.log
