#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c12; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_wide.py -q -m gpu -x > $O/wide.log 2>&1; echo "wide rc=$?" >> $O/rc.txt
timeout 600 python bench.py --model wide --batch 128 --steps 50 --warmup 5 --no-cpu-baseline --sites-json $O/sites_wide.json > $O/bench_wide.json 2> $O/bench_wide.err; echo "widebench rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -8 $O/wide.log; python - <<PY
import json
for f in ("bench_wide",):
    try:
        d=json.load(open("$O/%s.json"%f)); print(f, round(d['value']), d['ms_per_step'], round(d['e2e']['value']), d['step_roofline']['tensor_frac_of_sustained'], d['roofline']['kernel'], d['roofline']['frac'])
        for s in json.load(open("$O/sites_wide.json"))['sites']: print('   ', s)
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 $O/bench_wide.err
