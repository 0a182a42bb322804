#!/bin/bash
# Round 2, GPU pass N: threaded copy-out of large text blobs: GPU suite, then the default bench line.
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m "gpu and not slow" -x -q > $O/n_pytest.log 2>&1; tail -n 3 $O/n_pytest.log
timeout 600 python bench.py --no-cpu-baseline > $O/n_bench.json 2> $O/n_bench.err
python -c "import json; d=json.loads(open('$O/n_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['e2e']['ms_per_step'], d['ms_per_step'])"
