#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c6; mkdir -p $O
timeout 900 python -m pytest tests/test_gpu_cli.py -q -m gpu -x -s > $O/cli.log 2>&1; echo "cli rc=$?" >> $O/rc.txt
for S in 2 3 4; do
  echo "=== row stages $S" >> $O/probe.log
  B200_CONV_STAGES=$S timeout 300 python scratch/conv_probe.py 128 2>&1 | grep "variant 0" >> $O/probe.log
done
timeout 300 python cifar10cnn.py --job_name=worker --worker_hosts=localhost:2223 --max_steps=5000 --log_dir= > $O/cli_run.log 2>&1; echo "clirun rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -15 $O/cli.log; cat $O/probe.log; tail -3 $O/cli_run.log
