#!/bin/bash
# Two-GPU box: full GPU suite (multi-device tests not skipped), torchrun bench at N=2, P2TR bench at N=2, VG_TAP_GROUP A/B.
set -u
O=gpurun_out
nvidia-smi -L > $O/two_gpus.txt
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $O/pytest_two.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 > $O/bench_two.json 2> $O/bench_two.err
for v in "" _g32; do
  VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/tapgroup$v.json 2> $O/tapgroup$v.err
done
echo done
