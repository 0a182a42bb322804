#!/bin/bash
# Round 2, GPU pass I: the allocation-free entry point (alone in a fresh process, then after the whole suite as the driver runs it),
# smoke(), and the final bench lines of both arms.
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,driver_version --format=csv > $O/i_box.log 2>&1
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > $O/i_pytest_no_alloc_alone.log 2>&1; echo "alone rc=$?" >> $O/i_box.log
timeout 1800 python -m pytest tests -m gpu -x -q > $O/i_pytest_gpu_full.log 2>&1; echo "full rc=$?" >> $O/i_box.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/i_smoke.log 2>&1; echo "smoke rc=$?" >> $O/i_box.log
timeout 900 python bench.py > $O/i_bench.json 2> $O/i_bench.err; echo "bench rc=$?" >> $O/i_box.log
timeout 900 python bench.py --impl reference > $O/i_bench_reference.json 2> $O/i_bench_reference.err; echo "ref rc=$?" >> $O/i_box.log
cat $O/i_box.log; tail -n 3 $O/i_pytest_no_alloc_alone.log $O/i_pytest_gpu_full.log; cat $O/i_bench.json
