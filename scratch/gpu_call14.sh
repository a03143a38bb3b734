#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c14; mkdir -p $O
export B200_TRACE=1
python -c "
import sys; sys.path.insert(0,'.')
from b200cnn import build; build.build(force=True)" > $O/build.log 2>&1
for OV in 1 0; do
OVERLAP=$OV timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2954$OV scratch/trace_step_multi.py > $O/trace2_ov$OV.log 2>&1; echo "trace ov$OV rc=$?" >> $O/rc.txt
done
timeout 300 python scratch/trace_step.py 128 > $O/trace1.log 2>&1
cat $O/rc.txt; grep -v "^\*\|OMP\|^$" $O/trace2_ov1.log | tail -45; grep -v "^\*\|OMP\|^$" $O/trace2_ov0.log | tail -22
