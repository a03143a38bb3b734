#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c5; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_large_batch.py -q -m gpu -x -k "tma_strip" > $O/variants.log 2>&1; echo "variants rc=$?" >> $O/rc.txt
export B200_TRACE=1
python -c "
import sys; sys.path.insert(0,'.')
from b200cnn import build; build.build(force=True)" > $O/build.log 2>&1
for S in 0 11500; do
  echo "=== stagger $S" >> $O/probe.log
  B200_CONV_STAGGER_CYCLES=$S timeout 300 python scratch/conv_probe.py 4096 2>&1 | grep -A1 "variant 1" >> $O/probe.log
done
cat $O/rc.txt; tail -3 $O/variants.log; cat $O/probe.log
