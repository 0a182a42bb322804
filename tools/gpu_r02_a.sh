#!/bin/bash
# Round 2, GPU pass A: the GPU suite with the new tests, default bench, walk-kernel phase-split variants, and the first ncu captures
# the verdict asked for (walk kernel at round-2 state, Base58 DFA path `Cat$`, the P2TR tweak stage in its round-end configuration).
set -u
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > $O/a_smi.log 2>&1
nproc >> $O/a_smi.log
timeout 1500 python -m pytest tests -m "gpu and not slow" -x -q 2>&1 | tail -25 > $O/a_pytest.log
timeout 400 python bench.py > $O/a_bench.json 2> $O/a_bench.err
for v in "" _split4 _split2 _split1; do
  echo "== variant$v" >> $O/a_variants.log
  VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda$v.so timeout 300 python bench.py --steps 5 --warmup 3 --windows-per-step 16 --no-cpu-baseline 2>>$O/a_variants.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['e2e']['value'], d['clocks']['sm_mhz'], d['clocks'].get('power_w_max'), d['roofline']['launch_ms'])" >> $O/a_variants.log 2>&1
done
tools/refresh_ncu.sh capture r02_a > $O/a_refresh.log 2>&1
B="--steps 1 --warmup 3 --windows-per-step 2 --no-cpu-baseline"
timeout 300 python bench.py $B --format p2pkh --pattern 'Cat$' > $O/a_bench_b58.json 2> $O/a_bench_b58.err && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 8 -c 1 -o $O/ncu_walk_b58_r02_a -f python bench.py $B --format p2pkh --pattern 'Cat$' > $O/a_ncu_b58.log 2>&1
timeout 300 python bench.py $B --format p2tr --pattern '^bc1pcat' > $O/a_bench_p2tr.json 2> $O/a_bench_p2tr.err && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_taptweak -s 4 -c 1 -o $O/ncu_taptweak_r02_a -f python bench.py $B --format p2tr --pattern '^bc1pcat' > $O/a_ncu_tap.log 2>&1
./vgen_b200/peak_bench > $O/a_peak.json 2>&1
echo done
