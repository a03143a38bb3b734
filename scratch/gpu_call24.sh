#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c24; mkdir -p $O
for combo in "0 0" "14 74" "14 64" "16 64" "12 84" "10 96" "20 44" "26 64" "14 128"; do
  set -- $combo
  B200_WGRAD2_SPLITS=$1 B200_WGRAD1_CTAS=$2 timeout 200 python bench.py --steps 300 --warmup 20 --no-cpu-baseline 2> $O/err_$1_$2.log | tail -1 | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('splits $1 ctas $2:', round(d['ms_per_step']*1000,2), 'us cold,', round(128/d['value_hot_l2']*1e6,2), 'us hot (incl. 6.2 restore)')" >> $O/sweep.log
done
cat $O/sweep.log
