#!/bin/bash
# Round 2, GPU pass Q: ncu of the walk kernel with the Base58 suffix matcher (Cat$) at round end.
set -u
O=gpurun_out
mkdir -p $O
B="python bench.py --steps 1 --warmup 2 --windows-per-step 2 --no-cpu-baseline --format p2pkh --pattern Cat\$"
timeout 300 $B > $O/q_bench.json 2> $O/q_bench.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 6 -c 1 -o $O/ncu_walk_b58_r02_final -f $B > $O/q_ncu.log 2>&1
ls -la $O/ncu_walk_b58_r02_final.ncu-rep
