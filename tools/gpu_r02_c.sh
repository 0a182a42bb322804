#!/bin/bash
# Round 2, GPU pass C: GPU suite after the regex / text-blob / upload changes, default bench, a 2-"device" single-context bench on one GPU.
set -u
O=gpurun_out
mkdir -p $O
timeout 1700 python -m pytest tests -m "gpu and not slow" -q 2>&1 | tail -30 > $O/c_pytest.log
timeout 400 python bench.py > $O/c_bench.json 2> $O/c_bench.err
timeout 300 python bench.py --single-context --gpus 1 --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline > $O/c_single_1.json 2> $O/c_single_1.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > $O/c_ref.json 2> $O/c_ref.err
echo done
