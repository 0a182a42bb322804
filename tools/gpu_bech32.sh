#!/bin/bash
# Table-driven Bech32 checksum (match.cuh: vg_bech32_fill_table): GPU tests, the segwit DFA benches (C3 = P2WPKH dead$), and one
# full ncu capture of the P2WPKH walk kernel.
set -u
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $O/pytest_m.log
for cfg in "p2wpkh dead\$" "p2wpkh ^bc1qcat" "p2tr cat\$" "p2tr ^bc1pcat" "p2pkh ^1Cat"; do set -- $cfg; echo "== $1 $2"; timeout 200 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format $1 --pattern "$2" --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'hits/step', d['e2e']['hits_per_step'], d['config']['matcher'], 'clk', d['clocks']['sm_mhz'])"; done > $O/configs_m.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 8 -c 1 -o $O/ncu_walk_p2wpkh_m -f python bench.py --steps 1 --warmup 3 --windows-per-step 2 --format p2wpkh --pattern 'dead$' --no-cpu-baseline > $O/ncu_m.log 2>&1
echo done
