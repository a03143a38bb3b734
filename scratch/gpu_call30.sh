#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c30; mkdir -p $O
for m in write write+read; do
  B200_BENCH_FLUSH=$m timeout 300 python bench.py --steps 300 --warmup 20 --no-cpu-baseline 2> $O/err_$m.log | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('flush $m:', round(d['value']), round(d['ms_per_step']*1000,2), 'us cold;', round(d['value_hot_l2']), 'hot; e2e', round(d['e2e']['value']))" >> $O/out.log
  B200_WGRAD2_SPLITS=26 B200_WGRAD1_CTAS=128 B200_BENCH_FLUSH=$m timeout 300 python bench.py --steps 300 --warmup 20 --no-cpu-baseline 2>> $O/err_$m.log | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('flush $m (26,128):', round(d['value']), round(d['ms_per_step']*1000,2), 'us cold;', round(d['value_hot_l2']), 'hot; e2e', round(d['e2e']['value']))" >> $O/out.log
done
cat $O/out.log
