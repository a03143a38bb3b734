#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c17; mkdir -p $O
timeout 900 python scratch/bf16_errors.py $O/bf16_errors.json > $O/errors.log 2>&1; echo "errors rc=$?" >> $O/rc.txt
cat $O/rc.txt; cat $O/errors.log | tail -12
