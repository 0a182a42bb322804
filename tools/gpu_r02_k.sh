#!/bin/bash
# Round 2, GPU pass K: faster host encoders (confirm workers): dense-pattern end-to-end rates, same set as pass B (b_configs.log),
# then the BASELINE config shapes (tools/configs.sh).
set -u
O=gpurun_out
mkdir -p $O
for cfg in "p2pkh Cat\$" "p2pkh at\$" "p2sh-p2wpkh Zz\$" "p2pkh ^1C.*z\$" "ethereum ^0xAb" "p2pkh ^1Cat" "p2wpkh dead\$" "ethereum ^0xdead" "p2wpkh ^bc1qaa" "p2tr ^bc1pca"; do set -- $cfg; echo "== $1 $2"; timeout 200 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format $1 --pattern "$2" --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'hits/step', d['e2e']['hits_per_step'], d['geometry']['matcher'])"; done > $O/k_configs.log 2>&1
cat $O/k_configs.log
bash tools/configs.sh > $O/configs_r02_final.log 2>&1
cat $O/configs_r02_final.log
timeout 600 python -m pytest tests -m "gpu and not slow" -x -q > $O/k_pytest.log 2>&1; tail -n 3 $O/k_pytest.log
