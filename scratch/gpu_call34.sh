#!/bin/bash
# 2-GPU box: the multi-GPU pytest file (torchrun check + CLI 1 ps + 2 workers twice) on the push-exchange build
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c34; mkdir -p $O
export B200_P2P_TIMEOUT_S=10
timeout 400 python -m pytest tests/test_gpu_multi.py -q -m gpu -x > $O/multi.log 2>&1; echo "multi rc=$?" >> $O/rc.txt
tail -8 $O/multi.log; cat $O/rc.txt
