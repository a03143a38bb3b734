#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c26; mkdir -p $O
export B200_TRACE=1
python -c "
import sys; sys.path.insert(0,'.')
from b200cnn import build; build.build(force=True)" > $O/build.log 2>&1
timeout 300 python scratch/trace_step.py 128 > $O/trace_base.log 2>&1
B200_WGRAD2_SPLITS=14 B200_WGRAD1_CTAS=64 timeout 300 python scratch/trace_step.py 128 > $O/trace_co.log 2>&1
B200_WGRAD2_SPLITS=14 B200_WGRAD1_CTAS=128 timeout 300 python scratch/trace_step.py 128 > $O/trace_co128.log 2>&1
for f in trace_base trace_co trace_co128; do echo "== $f"; grep -A16 "^cold" $O/$f.log | grep "cold\|conv2_dgrad\|conv2_wgrad\|conv1_wgrad\|wgrad_reduce\|sgd"; grep -A16 "^hot" $O/$f.log | grep "hot\|conv2_wgrad\|conv1_wgrad\|wgrad_reduce\|sgd"; grep "wgrad CTA0" $O/$f.log; done
