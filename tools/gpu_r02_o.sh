#!/bin/bash
# Round 2, GPU pass O: P2TR fast path with an XYZZ accumulator (8M + 2S per mixed addition) against the Jacobian one
# (-DVG_TAP_XYZZ=0): parity, A/B on the same box, then an ncu capture of the stage.
set -u
O=gpurun_out
mkdir -p $O
timeout 900 python -m pytest tests -m "gpu and not slow" -x -q -k "p2tr or taproot or tap or digest or dispatch or published" > $O/o_pytest_p2tr.log 2>&1; tail -n 3 $O/o_pytest_p2tr.log
for v in "" _xyzz0 "" _xyzz0; do
  echo "== libvgen_cuda$v.so p2tr ^bc1pcat" >> $O/o_variants.log
  VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 8 --no-cpu-baseline --format p2tr --pattern '^bc1pcat' 2>>$O/o_variants.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'), d['e2e']['hits_per_step'])" >> $O/o_variants.log 2>&1
done
cat $O/o_variants.log
B="python bench.py --steps 1 --warmup 2 --windows-per-step 2 --no-cpu-baseline --format p2tr --pattern ^bc1pcat"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_taptweak -s 4 -c 1 -o $O/ncu_taptweak_r02_final -f $B > $O/o_ncu.log 2>&1
ls -la $O/ncu_taptweak_r02_final.ncu-rep
