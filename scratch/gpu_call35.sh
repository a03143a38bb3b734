#!/bin/bash
# 1 GPU: L2 prefetch of the backward kernels' late-read inputs (B200_L2_PREFETCH bit 0 = pool2/arg2, bit 1 = pool1/arg1/records)
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c35; mkdir -p $O
for pf in 3 0 1 2; do
B200_L2_PREFETCH=$pf timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_pf$pf.json 2> $O/bench_pf$pf.err; echo "bench pf$pf rc=$?" >> $O/rc.txt
done
for pf in 3 0; do
B200CNN_LIB=$PWD/scratch/libb200cnn_trace.so B200_L2_PREFETCH=$pf timeout 120 python scratch/trace_step.py 128 > $O/trace_pf$pf.log 2>&1; echo "trace pf$pf rc=$?" >> $O/rc.txt
done
timeout 300 python -m pytest tests/test_gpu_tensorcore.py -q -m gpu -x -k "step or determin" > $O/tests.log 2>&1; echo "tests rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -2 $O/tests.log
for pf in 3 0 1 2; do tail -1 $O/bench_pf$pf.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('pf$pf', round(d['value']), round(d['ms_per_step'],4), 'hot', round(d['value_hot_l2']), 'e2e', round(d['e2e']['value']))" 2>&1 | tail -1; done
for pf in 3 0; do echo "== trace pf$pf"; grep -A15 "^cold" $O/trace_pf$pf.log | head -16; grep "^hot" $O/trace_pf$pf.log; done
