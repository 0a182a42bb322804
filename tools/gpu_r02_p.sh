#!/bin/bash
# Round 2, GPU pass P: P2TR translation unit with the split reduction (fe.cuh VG_REDUCE_SPLIT) against the plain one: parity, A/B.
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m "gpu and not slow" -x -q -k "p2tr or taproot or tap or digest or dispatch or published" > $O/p_pytest_p2tr.log 2>&1; tail -n 3 $O/p_pytest_p2tr.log
for v in "" _rs0 "" _rs0; do
  echo "== libvgen_cuda$v.so p2tr ^bc1pcat" >> $O/p_variants.log
  VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 8 --no-cpu-baseline --format p2tr --pattern '^bc1pcat' 2>>$O/p_variants.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'), d['e2e']['hits_per_step'])" >> $O/p_variants.log 2>&1
done
cat $O/p_variants.log
