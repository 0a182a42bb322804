#!/bin/bash
# Round 2, GPU pass W: signed geometries of the P2TR table at the memory of the plain ones: 16 bits (64 MiB) against s17.1 (15
# windows, 67 MB); 24 bits (11 GiB, 11 windows) and 26 bits (43 GB, 10 windows) against s25.6 (10 windows, 17 GB).
set -u
O=gpurun_out
mkdir -p $O
for g in 16 s17.1 24 s25.6 26 16 s17.1 24 s25.6 26; do
  echo "== VGEN_TAP_BITS=$g p2tr ^bc1pcat" >> $O/w_variants.log
  VGEN_TRACE=1 VGEN_TAP_BITS=$g timeout 400 python bench.py --steps 3 --warmup 3 --windows-per-step 8 --no-cpu-baseline --format p2tr --pattern '^bc1pcat' 2>$O/w_variants_$g.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'), d['e2e']['hits_per_step'])" >> $O/w_variants.log 2>&1
  grep "tap table" $O/w_variants_$g.err | head -1 >> $O/w_variants.log
done
cat $O/w_variants.log
