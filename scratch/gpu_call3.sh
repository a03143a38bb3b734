#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c3; mkdir -p $O
export B200_TRACE=1
python -c "
import sys; sys.path.insert(0,'.')
from b200cnn import build; build.build(force=True)" > $O/build.log 2>&1
timeout 300 python scratch/conv_probe.py 4096 > $O/probe4096.log 2>&1; echo "probe rc=$?" >> $O/rc.txt
timeout 300 python scratch/conv_probe.py 128 > $O/probe128.log 2>&1; echo "probe128 rc=$?" >> $O/rc.txt
timeout 300 python scratch/trace_step.py 4096 > $O/trace4096.log 2>&1; echo "trace rc=$?" >> $O/rc.txt
timeout 300 python scratch/trace_step.py 128 > $O/trace128.log 2>&1; echo "trace128 rc=$?" >> $O/rc.txt
cat $O/rc.txt; cat $O/probe4096.log $O/probe128.log; cat $O/trace4096.log
