#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c36; mkdir -p $O
B200CNN_LIB=$PWD/scratch/libb200cnn_trace.so B200_L2_PREFETCH=0 timeout 120 python scratch/trace_step.py 128 12 > $O/trace.log 2>&1; echo "trace rc=$?"
grep "\[hot\]\|\[cold\]\|^hot\|^cold\|conv1_wgrad\|CTAs\|starts\|durations" $O/trace.log
