#!/bin/bash
# Table-driven Base58 radix conversion (match.cuh: vg_b58_fill_table): GPU tests, headline + Base58 DFA benches.
set -u
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $O/pytest_n2.log
for cfg in "p2pkh ^1Cat" "p2pkh Cat\$" "p2pkh (?i)satoshi" "p2sh-p2wpkh (?i)cat\$" "p2pkh-uncompressed Cat\$" "p2wpkh dead\$" "p2tr cat\$"; do set -- $cfg; echo "== $1 $2"; timeout 200 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format $1 --pattern "$2" --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'hits/step', d['e2e']['hits_per_step'], d['config']['matcher'], 'clk', d['clocks']['sm_mhz'])"; done > $O/configs_n2.log 2>&1
timeout 300 python bench.py > $O/bench_n2.json 2> $O/bench_n2.err
echo done
