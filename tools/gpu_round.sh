#!/bin/bash
# One GPU-box pass: full GPU test suite, default bench + reference arm, full-size C2 (incl. the alternation pattern),
# P2TR bench and a fresh ncu capture of the tweak stage. Outputs under gpurun_out/.
set -u
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $O/pytest_h.log
timeout 300 python bench.py > $O/bench_h.json 2> $O/bench_h.err
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_h_ref.json 2> $O/bench_h_ref.err
timeout 300 python tests/full_size_configs.py c2 > $O/full_c2_h.log 2>&1
timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/bench_h_p2tr.json 2> $O/bench_h_p2tr.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_taptweak -c 1 -o $O/ncu_taptweak_h -f python bench.py --steps 1 --warmup 3 --windows-per-step 1 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/ncu_tap_h.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_p2tr_h.csv python bench.py --steps 1 --warmup 3 --windows-per-step 2 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/ncu_tap_l_h.log 2>&1
echo done
