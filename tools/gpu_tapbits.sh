#!/bin/bash
# P2TR fixed-base table width A/B (VGEN_TAP_BITS), parity at 24 bits, scan phase trace, ncu of the Ethereum walk kernel.
set -u
O=gpurun_out
for b in 16 20 22 24; do
  VGEN_TAP_BITS=$b timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/tapbits_$b.json 2> $O/tapbits_$b.err
done
VGEN_TAP_BITS=24 timeout 600 python -m pytest tests -m gpu -x -q -k "p2tr or P2TR or taproot" 2>&1 | tail -5 > $O/pytest_tap24.log
VGEN_TRACE=1 timeout 300 python tools/e2e_breakdown.py > $O/e2e_trace.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -c 1 -o $O/ncu_walk_eth -f python bench.py --steps 1 --warmup 3 --windows-per-step 1 --format ethereum --pattern '^0xdead' --no-cpu-baseline > $O/ncu_eth.log 2>&1
echo done
