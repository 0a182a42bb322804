#!/bin/bash
# GPU tests, the default bench, then the ncu launch list and one full capture of the walk kernel for the same command.
set -u
O=gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > $O/pytest_k.log
timeout 300 python bench.py > $O/bench_k.json 2> $O/bench_k.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_k.csv python bench.py --steps 2 --warmup 3 --windows-per-step 4 --no-cpu-baseline > $O/ncu_l.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 8 -c 1 -o $O/ncu_walk_k -f python bench.py --steps 1 --warmup 3 --windows-per-step 2 --no-cpu-baseline > $O/ncu_w.log 2>&1
for cfg in "p2wpkh dead\$" "ethereum ^0xdead" "p2pkh ^1CatDog" "p2sh-p2wpkh ^3Cat" "p2pkh-uncompressed ^1Cat" "p2tr ^bc1pcat"; do set -- $cfg; echo "== $1 $2"; timeout 200 python bench.py --steps 3 --warmup 3 --windows-per-step 4 --format $1 --pattern "$2" --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'hits/step', d['e2e']['hits_per_step'], d['config']['matcher'])"; done > $O/configs_k.log 2>&1
echo done
