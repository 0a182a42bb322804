#!/bin/bash
# One GPU-box pass: full GPU test suite, default bench + reference arm, Ethereum (intervals / folded-DFA prefilter) and P2TR benches.
set -u
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > $O/pytest_n.log
timeout 300 python bench.py > $O/bench_n.json 2> $O/bench_n.err
timeout 300 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format ethereum --pattern '^0xdead' --no-cpu-baseline > $O/bench_n_eth.json 2> $O/bench_n_eth.err
timeout 300 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format ethereum --pattern 'F00$' --no-cpu-baseline > $O/bench_n_eth_dfa.json 2> $O/bench_n_eth_dfa.err
timeout 300 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format ethereum --pattern 'dead$' --no-cpu-baseline > $O/bench_n_eth_dfa2.json 2> $O/bench_n_eth_dfa2.err
timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/bench_n_p2tr.json 2> $O/bench_n_p2tr.err
echo done
