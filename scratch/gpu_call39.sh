#!/bin/bash
# 1 GPU: conv1 filter gradient with the band-wise scatter; wave splits
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c39; mkdir -p $O
timeout 120 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_wide.py -q -m gpu -x -k "conv1 or step or determin or wide" > $O/tests.log 2>&1; echo "tests rc=$?" >> $O/rc.txt
tail -12 $O/tests.log
if grep -q "tests rc=0" $O/rc.txt; then
timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_a.json 2> $O/bench_a.err; echo "bench a rc=$?" >> $O/rc.txt
B200_WGRAD1_CTAS=43 B200_WGRAD2_SPLITS=21 timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_c.json 2> $O/bench_c.err; echo "bench c rc=$?" >> $O/rc.txt
B200_WGRAD1_CTAS=64 B200_WGRAD2_SPLITS=14 timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_d.json 2> $O/bench_d.err; echo "bench d rc=$?" >> $O/rc.txt
fi
cat $O/rc.txt
for f in a c d; do tail -1 $O/bench_$f.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), 'hot', round(d['value_hot_l2']), 'e2e', round(d['e2e']['value'])); print([ (s['site'], round(s['us'],1)) for s in d['sites'] if 'wgrad' in s['site'] or 'dgrad' in s['site']])" 2>&1 | tail -2; done
