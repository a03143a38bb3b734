#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c8; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -q -m gpu -x > $O/tc.log 2>&1; echo "tc rc=$?" >> $O/rc.txt
timeout 400 python bench.py --batch 128 --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_b128.json 2> $O/bench_b128.err; echo "bench rc=$?" >> $O/rc.txt
export B200_TRACE=1
python -c "
import sys; sys.path.insert(0,'.')
from b200cnn import build; build.build(force=True)" > $O/build.log 2>&1
timeout 300 python scratch/trace_step.py 128 > $O/trace128.log 2>&1; echo "trace128 rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -3 $O/tc.log; python -c "
import json; d=json.load(open('$O/bench_b128.json')); print(d['value'], d['ms_per_step'], d['value_hot_l2'], d['e2e']['value'])"
cat $O/trace128.log
