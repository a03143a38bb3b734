#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c21; mkdir -p $O
timeout 300 python scratch/sanitize_step.py > $O/plain.log 2>&1; echo "plain rc=$?" >> $O/rc.txt
timeout 900 compute-sanitizer --tool memcheck --print-limit 20 python scratch/sanitize_step.py > $O/memcheck.log 2>&1; echo "memcheck rc=$?" >> $O/rc.txt
timeout 600 python -m pytest tests/test_gpu_tensorcore.py -q -m gpu -k "32x32 or trajectory" > $O/tc.log 2>&1; echo "tc rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -12 $O/memcheck.log; tail -3 $O/tc.log
