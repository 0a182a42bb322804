#!/bin/bash
# Round 2, GPU pass G: more resident warps (CTA size 288 / 320 at <= 113 / 102 registers), default and pipelined walk kernels.
set -u
O=gpurun_out
mkdir -p $O
for v in "" _b288 _b320 _pipe _pipe_b288 _pipe_b320; do
  echo "== variant$v" >> $O/g_variants.log
  VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 300 python bench.py --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline 2>>$O/g_variants.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'), d['roofline']['launch_ms'], d['geometry']['keys_per_window'])" >> $O/g_variants.log 2>&1
done
echo done
