#!/bin/bash
# Round 2, GPU pass X: P2TR stage 2 launch shape after this round's changes: 128-thread CTAs x 4 per SM (_tb128) and no per-key
# CTA barrier (_ts0) against the default (256 x 2, barrier per key); s25.6 table (short build).
set -u
O=gpurun_out
mkdir -p $O
for v in ${VARIANTS:-"" _tb128 _ts0 "" _tb128 _ts0}; do
  echo "== libvgen_cuda$v.so VGEN_TAP_BITS=s25.6" >> $O/x_variants.log
  VGEN_TAP_BITS=s25.6 VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda${v#def}.so timeout 400 python bench.py --steps 3 --warmup 3 --windows-per-step 8 --no-cpu-baseline --format p2tr --pattern '^bc1pcat' 2>>$O/x_variants.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'), d['e2e']['hits_per_step'])" >> $O/x_variants.log 2>&1
done
cat $O/x_variants.log
