#!/bin/bash
# Round 2, GPU pass B: the whole GPU suite (no -x), Base58 residue prefilter A/B, sanitizers.
set -u
O=gpurun_out
mkdir -p $O
timeout 1700 python -m pytest tests -m "gpu and not slow" -q 2>&1 | tail -40 > $O/b_pytest.log
for cfg in "p2pkh Cat\$" "p2pkh at\$" "p2sh-p2wpkh Zz\$" "p2pkh ^1C.*z\$" "ethereum ^0xAb" "p2pkh ^1Cat" "p2wpkh dead\$" "ethereum ^0xdead"; do set -- $cfg; echo "== $1 $2"; timeout 200 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format $1 --pattern "$2" --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'hits/step', d['e2e']['hits_per_step'], d['geometry']['matcher'])"; done > $O/b_configs.log 2>&1
bash tools/sanitize.sh r02 > $O/b_sanitize.log 2>&1
echo done
