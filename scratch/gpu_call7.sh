#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c7; mkdir -p $O
timeout 1200 python -m pytest tests -q -m gpu -x --durations=8 > $O/suite.log 2>&1; echo "suite rc=$?" >> $O/rc.txt
timeout 400 python bench.py --batch 128 --steps 300 --warmup 20 --no-cpu-baseline --sites-json $O/sites_b128.json > $O/bench_b128.json 2> $O/bench_b128.err; echo "bench rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -25 $O/suite.log; cat $O/bench_b128.json
