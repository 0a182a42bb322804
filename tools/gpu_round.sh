#!/bin/bash
# One GPU-box pass: full GPU test suite, default bench, Ethereum (32-bit-halves Keccak) and P2TR (auto table width) benches,
# ncu capture of the Ethereum walk kernel. Outputs under gpurun_out/.
set -u
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $O/pytest_j.log
timeout 300 python bench.py > $O/bench_j.json 2> $O/bench_j.err
timeout 300 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format ethereum --pattern '^0xdead' --no-cpu-baseline > $O/bench_j_eth.json 2> $O/bench_j_eth.err
timeout 300 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format ethereum --pattern 'F00$' --no-cpu-baseline > $O/bench_j_eth_dfa.json 2> $O/bench_j_eth_dfa.err
VGEN_TRACE=1 timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/bench_j_p2tr.json 2> $O/bench_j_p2tr.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -c 1 -o $O/ncu_walk_eth_b -f python bench.py --steps 1 --warmup 3 --windows-per-step 1 --format ethereum --pattern '^0xdead' --no-cpu-baseline > $O/ncu_eth_b.log 2>&1
echo done
