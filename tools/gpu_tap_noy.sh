#!/bin/bash
# P2TR stage 2: Y3 left out of the top-window addition (VG_TAP_LAST_NO_Y) A/B (lib_y = off), then the GPU tests with the default build.
set -u
O=gpurun_out
for v in "" _y; do for cfg in "p2tr ^bc1pcat" "p2tr cat\$"; do set -- $cfg; echo "== lib$v $1 $2"; VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 200 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format $1 --pattern "$2" --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'hits/step', d['e2e']['hits_per_step'], d['config']['matcher'], 'clk', d['clocks']['sm_mhz'])"; done; done > $O/tap_noy.log 2>&1
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $O/pytest_s.log
echo done
