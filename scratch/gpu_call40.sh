#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c40; mkdir -p $O
B200CNN_LIB=$PWD/scratch/libb200cnn_trace.so timeout 120 python scratch/trace_step.py 128 > $O/trace.log 2>&1; echo "trace rc=$?"
grep "\[hot\]\|^hot\|conv1_wgrad\|conv2_wgrad\|wgrad_reduce" $O/trace.log | head -12
