#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c20; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_tensorcore.py tests/test_gpu_parity.py -q -m gpu > $O/tc.log 2>&1; echo "tc rc=$?" >> $O/rc.txt
tail -4 $O/tc.log
python scratch/ncu_step.py > $O/plain_step2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"conv5x5_tc_kernel|conv1_wgrad_kernel|conv5x5_wgrad_tc_kernel|conv1_pool_fwd_kernel|fc_tail_rows_kernel|reduce_partials_tc" -s 7 -c 7 -o $O/r2_step_full python scratch/ncu_step.py > $O/ncu_full.log 2>&1; echo "ncu full rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -3 $O/ncu_full.log; ls -la $O
