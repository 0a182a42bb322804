#!/bin/bash
# Round-end evidence for HEAD: default bench (both arms), ncu launch list of the same command, full captures of the P2PKH and
# Ethereum walk kernels, the other configs.
set -u
O=gpurun_out
timeout 300 python bench.py > $O/bench_q.json 2> $O/bench_q.err
timeout 200 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_q_ref.json 2> $O/bench_q_ref.err
for cfg in "p2wpkh dead\$" "ethereum ^0xdead" "ethereum F00\$" "p2pkh ^1CatDog" "p2sh-p2wpkh ^3Cat" "p2pkh-uncompressed ^1Cat" "p2pkh Cat\$" "p2tr ^bc1pcat"; do set -- $cfg; echo "== $1 $2"; timeout 200 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format $1 --pattern "$2" --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'hits/step', d['e2e']['hits_per_step'], d['config']['matcher'], 'clk', d['clocks']['sm_mhz'])"; done > $O/configs_q.log 2>&1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_q.csv python bench.py --steps 2 --warmup 3 --windows-per-step 4 --no-cpu-baseline > $O/ncu_ql.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 8 -c 1 -o $O/ncu_walk_q -f python bench.py --steps 1 --warmup 3 --windows-per-step 2 --no-cpu-baseline > $O/ncu_qw.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 8 -c 1 -o $O/ncu_walk_eth_q -f python bench.py --steps 1 --warmup 3 --windows-per-step 2 --format ethereum --pattern '^0xdead' --no-cpu-baseline > $O/ncu_qe.log 2>&1
echo done
