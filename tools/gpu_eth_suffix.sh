#!/bin/bash
# Ethereum literal hex suffixes: last-digest-bits prefilter (vg_dfa_dev::chk_mask in vg_match_eth). GPU tests, then the Ethereum benches.
set -u
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $O/pytest_u.log
for cfg in "ethereum dead\$" "ethereum F00\$" "ethereum 1234\$" "ethereum ^0xdead"; do set -- $cfg; echo "== $1 $2"; timeout 200 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format $1 --pattern "$2" --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'hits/step', d['e2e']['hits_per_step'], d['config']['matcher'], 'clk', d['clocks']['sm_mhz'])"; done > $O/eth_suffix.log 2>&1
echo done
