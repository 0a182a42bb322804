#!/bin/bash
# Round 2, GPU pass R: signed mixed-radix geometries of the P2TR fixed-base table: small ones against the oracle, then the
# 9-window 111.7 GB table (s28.4) against the 26-bit one on the same box, and P2TR parity tests with the big table forced.
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m "gpu and not slow" -x -q -k "p2tr or taproot or tap or digest or dispatch or published" > $O/r_pytest_p2tr.log 2>&1; tail -n 3 $O/r_pytest_p2tr.log
for g in 26 s28.4 26 s28.4; do
  echo "== VGEN_TAP_BITS=$g p2tr ^bc1pcat" >> $O/r_variants.log
  VGEN_TRACE=1 VGEN_TAP_BITS=$g timeout 400 python bench.py --steps 3 --warmup 3 --windows-per-step 8 --no-cpu-baseline --format p2tr --pattern '^bc1pcat' 2>$O/r_variants_$g.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'), d['e2e']['hits_per_step'])" >> $O/r_variants.log 2>&1
  grep "tap table" $O/r_variants_$g.err | head -2 >> $O/r_variants.log
done
cat $O/r_variants.log
VGEN_TAP_BITS=s28.4 timeout 900 python -m pytest tests -m "gpu and not slow" -x -q -k "p2tr and not geometries and not long_job" > $O/r_pytest_big.log 2>&1; tail -n 3 $O/r_pytest_big.log
nvidia-smi --query-gpu=memory.total,memory.used --format=csv
