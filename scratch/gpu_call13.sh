#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c13; mkdir -p $O
export B200_TRACE=1
python -c "
import sys; sys.path.insert(0,'.')
from b200cnn import build; build.build(force=True)" > $O/build.log 2>&1
timeout 300 python scratch/wide_probe.py 128 > $O/probe.log 2>&1; echo "probe rc=$?" >> $O/rc.txt
cat $O/rc.txt; cat $O/probe.log | tail -20
