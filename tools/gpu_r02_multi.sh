#!/bin/bash
# Round 2, multi-GPU pass (N = number of GPUs on the box): single-context path vs one-process-per-GPU, multi-device tests,
# fault injection, full-size BASELINE configs in one context, C5 strong-scaling point.
set -u
N=${1:-8}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=index,name --format=csv > $O/m${N}_smi.log 2>&1; nproc >> $O/m${N}_smi.log
timeout 900 python -m pytest tests/test_gpu_parity.py -q -k "multi_device or failed_device or sharded" 2>&1 | tail -15 > $O/m${N}_pytest_multi.log
for n in $N $((N/2)) 1; do
  [ $n -ge 1 ] || continue
  timeout 300 python bench.py --single-context --gpus $n --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline > $O/m${N}_single_$n.json 2> $O/m${N}_single_$n.err
done
timeout 400 python bench.py --gpus $N --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline > $O/m${N}_torchrun_$N.json 2> $O/m${N}_torchrun_$N.err
timeout 600 python bench.py --single-context --gpus $N --c5 > $O/m${N}_c5_$N.json 2> $O/m${N}_c5_$N.err
timeout 900 python -m pytest tests/test_full_size_configs.py -q -s 2>&1 | tail -15 > $O/m${N}_pytest_full.log
echo done
