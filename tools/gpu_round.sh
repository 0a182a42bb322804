#!/bin/bash
# One GPU-box pass: full GPU test suite, default bench, format variants, P2TR profile. Outputs under gpurun_out/.
set -u
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $O/pytest_e.log
timeout 300 python bench.py > $O/bench_e.json 2> $O/bench_e.err
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_e_ref.json 2> $O/bench_e_ref.err
for v in "" _ku; do
  VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 300 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format ethereum --pattern '^0xdead' --no-cpu-baseline > $O/bench_e_eth$v.json 2> $O/bench_e_eth$v.err
done
timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/bench_e_p2tr.json 2> $O/bench_e_p2tr.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_taptweak -c 1 -o $O/ncu_taptweak -f python bench.py --steps 1 --warmup 3 --windows-per-step 1 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/ncu_tap.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_p2tr.csv python bench.py --steps 1 --warmup 3 --windows-per-step 2 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/ncu_tap_l.log 2>&1
./vgen_b200/vgen range --puzzle 30 -p '^1CatD' -c 0 -o jsonl > $O/cli_p30.jsonl 2> $O/cli_p30.err
./vgen_b200/vgen estimate -p '^bc1pcat' -f p2tr > $O/cli_estimate.log 2>&1
echo done
