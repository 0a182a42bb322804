#!/bin/bash
# Round 2, GPU pass D: checked build (device asserts + buffer canaries) over every kernel, then the GPU suite on the product build.
set -u
O=gpurun_out
mkdir -p $O
bash tools/sanitize.sh r02 > $O/d_checked.log 2>&1
timeout 1700 python -m pytest tests -m "gpu and not slow" -q 2>&1 | tail -12 > $O/d_pytest.log
timeout 300 python bench.py --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline > $O/d_bench.json 2> $O/d_bench.err
echo done
