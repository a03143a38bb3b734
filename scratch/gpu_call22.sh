#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c22; mkdir -p $O
timeout 1200 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_large_batch.py -q -m gpu > $O/tc.log 2>&1; echo "tc rc=$?" >> $O/rc.txt
tail -4 $O/tc.log
for B in 128 512 1024 4096; do
  timeout 400 python bench.py --batch $B --steps 200 --warmup 10 --no-cpu-baseline --sites-json $O/r2_bf16_sites_b$B.json > $O/r2_bf16_bench_b$B.json 2> $O/bench_b$B.err; echo "bench$B rc=$?" >> $O/rc.txt
done
cat $O/rc.txt
for B in 128 512 1024 4096; do tail -1 $O/r2_bf16_bench_b$B.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print($B, round(d['value']), round(d['ms_per_step'],4), round(d.get('value_hot_l2',0)), round(d['e2e']['value']), d['step_roofline']['tensor_frac_of_sustained'], d['roofline']['kernel'], round(d['roofline']['frac'],4))"; done
