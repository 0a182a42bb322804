#!/bin/bash
# Round 2, GPU pass T: P2TR stage 2 with an L2 prefetch of the key's table entries against none (_pf0), 9-window signed table and
# 26-bit table; parity subset first.
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m "gpu and not slow" -x -q -k "p2tr or taproot or tap or digest or dispatch or published" > $O/t_pytest_p2tr.log 2>&1; tail -n 3 $O/t_pytest_p2tr.log
for g in s28.4 26; do for v in "" _pf0 "" _pf0; do
  echo "== libvgen_cuda$v.so VGEN_TAP_BITS=$g" >> $O/t_variants.log
  VGEN_TAP_BITS=$g VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 400 python bench.py --steps 3 --warmup 3 --windows-per-step 8 --no-cpu-baseline --format p2tr --pattern '^bc1pcat' 2>>$O/t_variants.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'), d['e2e']['hits_per_step'])" >> $O/t_variants.log 2>&1
done; done
cat $O/t_variants.log
