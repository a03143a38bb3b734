#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c11; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_wide.py tests/test_gpu_tensorcore.py -q -m gpu -x > $O/tc.log 2>&1; echo "tc rc=$?" >> $O/rc.txt
timeout 400 python bench.py --batch 128 --steps 300 --warmup 20 --no-cpu-baseline > $O/bench_b128.json 2> $O/bench_b128.err; echo "bench rc=$?" >> $O/rc.txt
timeout 600 python bench.py --model wide --batch 128 --steps 50 --warmup 5 --cpu-seconds 8 --sites-json $O/sites_wide.json > $O/bench_wide.json 2> $O/bench_wide.err; echo "wide rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -3 $O/tc.log; python - <<PY
import json
for f in ("bench_b128","bench_wide"):
    try:
        d=json.load(open("$O/%s.json"%f)); print(f, round(d['value']), d['ms_per_step'], round(d['e2e']['value']), d['step_roofline']['tensor_frac_of_sustained'], d['roofline']['kernel'], d['roofline']['frac'])
        for s in d['sites']: print('   ', s)
        print(d.get('cpu_baseline'))
    except Exception as e: print(f, 'ERR', e)
PY
tail -5 $O/bench_wide.err
