#!/bin/bash
# Two-GPU box: full GPU suite (multi-device tests not skipped), torchrun bench at N=2 for both arms.
set -u
O=gpurun_out
nvidia-smi -L > $O/two_gpus.txt
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $O/pytest_two.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > $O/bench_two.json 2> $O/bench_two.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > $O/bench_two_ref.json 2> $O/bench_two_ref.err
echo done
