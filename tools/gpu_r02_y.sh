#!/bin/bash
# Round 2, GPU pass Y (last commit): the whole GPU suite as the driver runs it, smoke(), the default bench line.
set -u
O=gpurun_out
mkdir -p $O
timeout 1800 python -m pytest tests -m gpu -x -q > $O/y_pytest_gpu_full.log 2>&1; tail -n 3 $O/y_pytest_gpu_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/y_smoke.log 2>&1; tail -n 1 $O/y_smoke.log
timeout 900 python bench.py --no-cpu-baseline > $O/y_bench.json 2> $O/y_bench.err; echo "bench rc=$?"
python -c "import json; d=json.loads(open('$O/y_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['roofline']['frac'])"
