#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c31; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_wide.py -q -m gpu -x > $O/wide.log 2>&1; echo "wide rc=$?" >> $O/rc.txt
tail -5 $O/wide.log
timeout 300 python bench.py --model wide --steps 100 --warmup 10 --no-cpu-baseline --sites-json $O/sites_wide.json > $O/bench_wide.json 2> $O/wide.err; echo "widebench rc=$?" >> $O/rc.txt
B200_CONV_WIDE_IPC=1 timeout 300 python bench.py --model wide --steps 100 --warmup 10 --no-cpu-baseline > $O/bench_wide_ipc1.json 2> $O/wide1.err; echo "widebench1 rc=$?" >> $O/rc.txt
cat $O/rc.txt
for f in bench_wide bench_wide_ipc1; do tail -1 $O/$f.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), round(d['value_hot_l2']), 'e2e', round(d['e2e']['value']), round(d['step_roofline']['tensor_frac_of_sustained'],4)); [print('   ', s) for s in d['sites'][:5]]"; done
tail -3 $O/wide.err
