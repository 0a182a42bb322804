#!/bin/bash
# Segwit literal-suffix patterns: checksum prefilter before the DFA (vg_dfa_dev::chk_mask). GPU tests, then C3 and P2TR benches.
set -u
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $O/pytest_t.log
for cfg in "p2wpkh dead\$" "p2wpkh de.d\$" "p2tr cat\$" "p2pkh ^1Cat"; do set -- $cfg; echo "== $1 $2"; timeout 200 python bench.py --steps 4 --warmup 3 --windows-per-step 8 --format $1 --pattern "$2" --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'hits/step', d['e2e']['hits_per_step'], d['config']['matcher'], 'clk', d['clocks']['sm_mhz'])"; done > $O/chk_suffix.log 2>&1
echo done
