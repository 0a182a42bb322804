#!/bin/bash
# Round 2, GPU pass S (final binary): the checked build over every kernel, the whole GPU suite as the driver runs it, smoke(),
# and both bench arms.
set -u
O=gpurun_out
mkdir -p $O
bash tools/sanitize.sh r02_s > $O/s_sanitize.log 2>&1; cat $O/s_sanitize.log
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/s_pytest_no_alloc_alone.log 2>&1; tail -n 2 $O/s_pytest_no_alloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > $O/s_pytest_gpu_full.log 2>&1; tail -n 3 $O/s_pytest_gpu_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/s_smoke.log 2>&1; tail -n 1 $O/s_smoke.log
timeout 900 python bench.py --impl reference > $O/s_bench_reference.json 2> $O/s_bench_reference.err; echo "ref rc=$?"
timeout 900 python bench.py > $O/s_bench.json 2> $O/s_bench.err; echo "bench rc=$?"
python -c "import json; d=json.loads(open('$O/s_bench.json').read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['roofline']['frac'], d['roofline']['binding_unit']['frac'], d['cpu_baseline']['value'])"
