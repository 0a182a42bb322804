#!/bin/bash
# Round 2, GPU pass F: single-hash-copy software-pipelined walk kernel against the default: parity subset, A/B bench, ncu.
set -u
O=gpurun_out
mkdir -p $O
VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda_pipe.so timeout 900 python -m pytest tests/test_gpu_parity.py -q -x -k "hash160_batch or generate_batch or config_c1 or config_c3 or scan_edges or multi_window or two_lanes or small" 2>&1 | tail -6 > $O/f_pytest_pipe.log
for v in "" _pipe "" _pipe; do
  for cfg in "p2pkh ^1Cat" "p2wpkh dead\$"; do set -- $cfg
  echo "== variant$v $1 $2" >> $O/f_variants.log
  VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 300 python bench.py --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline --format $1 --pattern "$2" 2>>$O/f_variants.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'), d['roofline']['launch_ms'])" >> $O/f_variants.log 2>&1
  done
done
B="--steps 1 --warmup 3 --windows-per-step 2 --no-cpu-baseline"
VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda_pipe.so timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 8 -c 1 -o $O/ncu_walk_pipe_r02_b -f python bench.py $B > $O/f_ncu_pipe.log 2>&1
echo done
