#!/bin/bash
# Round 2, GPU pass M (8-GPU box): one context over N GPUs against one process per GPU, after the host-encoder rewrite.
set -u
N=${1:-8}
O=gpurun_out
mkdir -p $O
timeout 300 python bench.py --single-context --gpus $N --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline > $O/m${N}_single.json 2> $O/m${N}_single.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline > $O/m${N}_torchrun.json 2> $O/m${N}_torchrun.err
for f in $O/m${N}_single.json $O/m${N}_torchrun.json; do python -c "import sys,json; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('$f', d['value'], d['e2e']['value'], d['clocks']['sm_mhz'])"; done
nproc
