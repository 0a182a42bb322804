#!/bin/bash
# P2TR after the grouped table build: parity tests (16- and 24-bit tables), build times (VGEN_TRACE), widths 24 and 26.
set -u
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "p2tr or P2TR or taproot" 2>&1 | tail -5 > $O/pytest_tap_k.log
VGEN_TAP_BITS=24 timeout 600 python -m pytest tests -m gpu -x -q -k "p2tr or P2TR or taproot" 2>&1 | tail -5 >> $O/pytest_tap_k.log
for b in 16 24 26; do
  VGEN_TRACE=1 VGEN_TAP_BITS=$b timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format p2tr --pattern '^bc1pcat' --no-cpu-baseline > $O/tapbits_k_$b.json 2> $O/tapbits_k_$b.err
done
echo done
