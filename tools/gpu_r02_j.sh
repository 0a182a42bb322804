#!/bin/bash
# Round 2, GPU pass J (round-end binary): checked build + canaries over every kernel incl. the allocation-free entry point,
# then the ncu launch list and the --set full capture of the walk kernel that bench.py's roofline quotes, then the per-format lines.
set -u
O=gpurun_out
mkdir -p $O
bash tools/sanitize.sh r02_final > $O/j_sanitize.log 2>&1
cat $O/j_sanitize.log
bash tools/refresh_ncu.sh capture r02_final > $O/j_refresh.log 2>&1
tail -2 $O/j_refresh.log
bash tools/configs.sh > $O/configs_r02_final.log 2>&1
timeout 300 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --no-cpu-baseline --format p2tr --pattern '^bc1pcat' 2>/dev/null | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('== P2TR ^bc1pcat'); print(d['value'], 'e2e', d['e2e']['value'], 'clk', d['clocks']['sm_mhz'], 'power', d['clocks'].get('power_w_max'))" >> $O/configs_r02_final.log 2>&1
cat $O/configs_r02_final.log
