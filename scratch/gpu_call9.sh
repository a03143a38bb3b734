#!/bin/bash
cd "$GRAFT_REPO_ROOT" || exit 1
O=gpurun_out/c9; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -q -m gpu -x -s > $O/multi.log 2>&1; echo "multi rc=$?" >> $O/rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 200 --warmup 10 > $O/bench_2gpu.json 2> $O/bench_2gpu.err; echo "bench2 rc=$?" >> $O/rc.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --steps 200 --warmup 10 --batch 1024 > $O/bench_2gpu_b1024.json 2> $O/bench_2gpu_b1024.err; echo "bench2b rc=$?" >> $O/rc.txt
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29535 bench.py --impl reference --gpus 2 --steps 60 --warmup 3 > $O/bench_ref2.json 2> $O/bench_ref2.err; echo "ref2 rc=$?" >> $O/rc.txt
cat $O/rc.txt; tail -30 $O/multi.log; cat $O/bench_2gpu.json | cut -c1-600; cat $O/bench_2gpu_b1024.json | cut -c1-300; cat $O/bench_ref2.json | cut -c1-300
