#!/bin/bash
# P2TR inversion group A/B (32 = default build, 64 = variant), parity with the new default, P2PKH DFA-path rates.
set -u
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "p2tr or P2TR or taproot or match_sets" 2>&1 | tail -5 > $O/pytest_tap_m.log
for v in "" _g64; do
  VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/tapgroup_m$v.json 2> $O/tapgroup_m$v.err
done
timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 8 --format p2pkh --pattern 'Cat$' --no-cpu-baseline > $O/bench_m_p2pkh_suffix.json 2> $O/bench_m_p2pkh_suffix.err
timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 8 --format p2pkh --pattern '(?i)satoshi' --no-cpu-baseline > $O/bench_m_p2pkh_sub.json 2> $O/bench_m_p2pkh_sub.err
echo done
