#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c10; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_wide.py -q -m gpu -x > $O/wide.log 2>&1; echo "wide rc=$?" >> $O/rc.txt
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_large_batch.py -q -m gpu -x -k "not 4096" > $O/tc.log 2>&1; echo "tc rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -30 $O/wide.log; tail -5 $O/tc.log
