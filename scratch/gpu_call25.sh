#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c25; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -k "host" > $O/host.log 2>&1; echo "host rc=$?" >> $O/rc.txt
tail -5 $O/host.log
timeout 300 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/rc.txt
B200_WGRAD2_SPLITS=14 B200_WGRAD1_CTAS=64 timeout 300 python bench.py --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_co.json 2> $O/bench_co.err; echo "bench_co rc=$?" >> $O/rc.txt
timeout 300 python bench.py --batch 1024 --steps 100 --warmup 10 --no-cpu-baseline > $O/bench_b1024.json 2> $O/bench_b1024.err; echo "bench1024 rc=$?" >> $O/rc.txt
cat $O/rc.txt
for f in bench bench_co bench_b1024; do tail -1 $O/$f.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), round(d['value_hot_l2']), 'e2e', round(d['e2e']['value']), d['e2e']['last_loss'])"; done
tail -3 $O/bench.err
