#!/bin/bash
# final single-GPU validation of the round-2 build: whole GPU suite, smoke, bench (+ large batch, wide, CLI loop), trace, ncu launch list
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c41; mkdir -p $O
timeout 600 python -m pytest tests -q -m gpu --durations=5 > $O/suite.log 2>&1; echo "suite rc=$?" >> $O/rc.txt
tail -4 $O/suite.log
python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/rc.txt
timeout 300 python bench.py --steps 300 --warmup 20 --sites-json $O/r2_bf16_sites.json > $O/r2_bf16_bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/rc.txt
for B in 512 1024 4096; do timeout 200 python bench.py --batch $B --steps 100 --warmup 10 --no-cpu-baseline > $O/r2_bf16_bench_b$B.json 2> $O/bench_b$B.err; echo "bench$B rc=$?" >> $O/rc.txt; done
timeout 200 python bench.py --model wide --steps 100 --warmup 10 --no-cpu-baseline > $O/r2_wide_bench.json 2> $O/wide.err; echo "wide rc=$?" >> $O/rc.txt
timeout 200 python cifar10cnn.py --job_name=worker --worker_hosts=localhost:2223 --max_steps=10000 --log_dir= > $O/r2_cli_loop.log 2>&1; echo "cli rc=$?" >> $O/rc.txt
B200CNN_LIB=$PWD/scratch/libb200cnn_trace.so timeout 120 python scratch/trace_step.py 128 > $O/trace.log 2>&1; echo "trace rc=$?" >> $O/rc.txt
if grep -q "bench rc=0" $O/rc.txt; then
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/ncu.log 2>&1; echo "ncu rc=$?" >> $O/rc.txt
fi
cat $O/rc.txt
for f in r2_bf16_bench r2_bf16_bench_b512 r2_bf16_bench_b1024 r2_bf16_bench_b4096 r2_wide_bench; do tail -1 $O/$f.json | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), round(d['value_hot_l2']), 'e2e', round(d['e2e']['value']), d['roofline']['kernel'], round(d['roofline']['frac'],3), d['gpu_launches'], (d.get('cpu_baseline') or {}).get('value'), round(d['step_roofline']['tensor_frac_of_sustained'],4))"; done
tail -1 $O/r2_cli_loop.log; tail -2 $O/smoke.log; grep "^hot\|^cold" $O/trace.log
