#!/bin/bash
# Round 2, scaling pass on an N-GPU box: single-context bench (one process, N devices), the torchrun arm for comparison, and the
# C5 strong-scaling point (fixed 2^40-key window on N GPUs).
set -u
N=${1:-1}
O=gpurun_out
mkdir -p $O
timeout 300 python bench.py --single-context --gpus $N --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline > $O/s${N}_single.json 2> $O/s${N}_single.err
if [ $N -gt 1 ]; then timeout 400 python bench.py --gpus $N --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline > $O/s${N}_torchrun.json 2> $O/s${N}_torchrun.err; fi
timeout 900 python bench.py --single-context --gpus $N --c5 > $O/s${N}_c5.json 2> $O/s${N}_c5.err
echo done
