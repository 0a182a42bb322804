#!/bin/bash
# Round 2, GPU pass U: the GPU suite on the last binary (after the table-entry refactor) and the P2TR line with the default table choice.
set -u
O=gpurun_out
mkdir -p $O
timeout 1200 python -m pytest tests -m "gpu and not slow" -x -q > $O/u_pytest.log 2>&1; tail -n 3 $O/u_pytest.log
VGEN_TRACE=1 timeout 400 python bench.py --steps 3 --warmup 3 --windows-per-step 8 --no-cpu-baseline --format p2tr --pattern '^bc1pcat' > $O/u_bench_p2tr.json 2> $O/u_bench_p2tr.err
python -c "import json; d=json.loads(open('$O/u_bench_p2tr.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['clocks']['sm_mhz'])"; grep "tap table" $O/u_bench_p2tr.err | head -3
