#!/bin/bash
# round-2 call 2: TMA strip + two-CTAs-per-SM conv2 kernels
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c2; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_large_batch.py -q -m gpu -x -k "tma_strip or multi_image" > $O/variants.log 2>&1; echo "variants rc=$?" >> $O/rc.txt
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_parity.py -q -m gpu -x > $O/suite.log 2>&1; echo "suite rc=$?" >> $O/rc.txt
timeout 900 python -m pytest tests/test_gpu_large_batch.py -q -m gpu -x -k "full_step or batch_4096" > $O/large.log 2>&1; echo "large rc=$?" >> $O/rc.txt
for B in 128 1024 4096; do
  timeout 400 python bench.py --batch $B --steps 200 --warmup 10 --no-cpu-baseline --sites-json $O/sites_b$B.json > $O/bench_b$B.json 2> $O/bench_b$B.err; echo "bench$B rc=$?" >> $O/rc.txt
done
cat $O/rc.txt
tail -5 $O/variants.log; tail -3 $O/suite.log; tail -3 $O/large.log
