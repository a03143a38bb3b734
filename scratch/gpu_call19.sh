#!/bin/bash
# final single-GPU measurements + ncu evidence of round 2
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c19; mkdir -p $O
timeout 1500 python -m pytest tests -q -m gpu --durations=5 > $O/suite.log 2>&1; echo "suite rc=$?" >> $O/rc.txt
tail -6 $O/suite.log
timeout 400 python bench.py --steps 300 --warmup 20 --sites-json $O/r2_bf16_sites.json > $O/r2_bf16_bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/rc.txt
for B in 512 1024 4096; do
  timeout 400 python bench.py --batch $B --steps 200 --warmup 10 --no-cpu-baseline --sites-json $O/r2_bf16_sites_b$B.json > $O/r2_bf16_bench_b$B.json 2> $O/bench_b$B.err; echo "bench$B rc=$?" >> $O/rc.txt
done
timeout 600 python bench.py --model wide --steps 100 --warmup 10 --cpu-seconds 8 --sites-json $O/r2_wide_sites.json > $O/r2_wide_bench.json 2> $O/bench_wide.err; echo "wide rc=$?" >> $O/rc.txt
timeout 400 python bench.py --mode fp32 --steps 50 --warmup 5 --no-cpu-baseline --sites-json $O/r2_fp32_sites.json > $O/r2_fp32_bench.json 2> $O/bench_fp32.err; echo "fp32 rc=$?" >> $O/rc.txt
timeout 300 python bench.py --impl reference --steps 200 --warmup 3 > $O/r2_reference_bench.json 2> $O/bench_ref.err; echo "ref rc=$?" >> $O/rc.txt
timeout 300 python cifar10cnn.py --job_name=worker --worker_hosts=localhost:2223 --max_steps=10000 --log_dir= > $O/r2_cli_loop.log 2>&1; echo "cli rc=$?" >> $O/rc.txt
# ncu: launch list, then --set full of the hot kernels (plain run first, as the recipe asks)
python scratch/ncu_step.py > $O/plain_step.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/r2_bf16_launches.csv python scratch/ncu_step.py > $O/ncu_launch.log 2>&1; echo "ncu launches rc=$?" >> $O/rc.txt
python scratch/ncu_step.py > $O/plain_step2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"conv5x5_tc_kernel|conv1_wgrad_kernel|conv5x5_wgrad_tc_kernel|conv1_pool_fwd_kernel|fc_tail_rows_kernel" -s 30 -c 12 -o $O/r2_step_full python scratch/ncu_step.py > $O/ncu_full.log 2>&1; echo "ncu full rc=$?" >> $O/rc.txt
python scratch/ncu_large.py > $O/plain_large.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"conv5x5_tc_kernel" -c 8 -o $O/r2_large_full python scratch/ncu_large.py > $O/ncu_large.log 2>&1; echo "ncu large rc=$?" >> $O/rc.txt
cat $O/rc.txt
for f in r2_bf16_bench r2_bf16_bench_b512 r2_bf16_bench_b1024 r2_bf16_bench_b4096 r2_wide_bench r2_fp32_bench r2_reference_bench; do tail -1 $O/$f.json | python -c "
import json,sys
try:
    d=json.loads(sys.stdin.read()); print('$f', round(d['value']), round(d['ms_per_step'],4), round(d.get('value_hot_l2',0)), round(d['e2e']['value']), d.get('step_roofline',{}).get('tensor_frac_of_sustained'), (d.get('roofline') or {}).get('kernel'), (d.get('roofline') or {}).get('frac'), (d.get('cpu_baseline') or {}).get('value'), (d.get('cpu_baseline_1ps_2workers') or {}).get('value'))
except Exception as e: print('$f ERR', e)"; done
tail -1 $O/r2_cli_loop.log
