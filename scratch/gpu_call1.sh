#!/bin/bash
# round-2 call 1: large-batch parity, whole GPU suite, bench sweep over the batch sizes of config 4
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c1; mkdir -p $O
nvidia-smi --query-gpu=name,memory.total --format=csv > $O/gpu.txt 2>&1
free -g > $O/host.txt; nproc >> $O/host.txt
timeout 900 python -m pytest tests/test_gpu_large_batch.py -q -m gpu -x --durations=10 > $O/large_batch.log 2>&1; echo "large rc=$?" >> $O/rc.txt
timeout 900 python -m pytest tests -q -m gpu --deselect tests/test_gpu_large_batch.py --durations=5 > $O/suite.log 2>&1; echo "suite rc=$?" >> $O/rc.txt
for B in 128 512 1024 4096; do
  extra="--no-cpu-baseline"; [ $B = 128 ] && extra=""
  timeout 400 python bench.py --batch $B --steps 200 --warmup 10 $extra --sites-json $O/sites_b$B.json > $O/bench_b$B.json 2> $O/bench_b$B.err; echo "bench$B rc=$?" >> $O/rc.txt
done
timeout 300 python bench.py --impl reference --steps 100 --warmup 3 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?" >> $O/rc.txt
cat $O/rc.txt
tail -5 $O/large_batch.log
tail -3 $O/suite.log
