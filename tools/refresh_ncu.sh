#!/bin/bash
# Regenerates the ncu evidence bench.py quotes (profiles/walk_kernel_ncu.json: executed instructions per key, ALU-pipe
# instructions per key, DRAM bytes per launch of the walk kernel) from the CURRENT binary.
#
#   on the GPU box (through gpurun):   tools/refresh_ncu.sh capture  <tag>     -> gpurun_out/ncu_walk_<tag>.ncu-rep, launches_<tag>.csv
#   here, after the call returned:      tools/refresh_ncu.sh summarise <tag>    -> profiles/ncu_walk_<tag>_summary.txt, profiles/launches_<tag>.csv,
#                                                                                profiles/walk_kernel_ncu.json (with commit, date, command)
# The capture only runs after the same command has exited 0 without ncu (B200_PROFILING.md).
set -u
cmd=${1:?capture|summarise}; tag=${2:?tag}
O=gpurun_out
BENCH="python bench.py --steps 1 --warmup 3 --windows-per-step 2 --no-cpu-baseline"
case "$cmd" in
capture)
    mkdir -p $O
    timeout 300 $BENCH > $O/bench_for_ncu_$tag.json 2> $O/bench_for_ncu_$tag.err || { echo "bench failed, no capture"; exit 1; }
    timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_$tag.csv $BENCH > $O/ncu_launches_$tag.log 2>&1
    timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_walk -s 8 -c 1 -o $O/ncu_walk_$tag -f $BENCH > $O/ncu_full_$tag.log 2>&1
    ls -la $O/ncu_walk_$tag.ncu-rep
    ;;
summarise)
    keys=$(python -c "import json;print(json.load(open('$O/bench_for_ncu_$tag.json'))['geometry']['keys_per_window'])")
    python tools/ncu_summary.py $O/ncu_walk_$tag.ncu-rep --kernel k_walk --keys $keys --out profiles/ncu_walk_${tag}_summary.txt \
        --json profiles/walk_kernel_ncu.json --command "$BENCH" --note "k_walk<0,0> (P2PKH, intervals ^1Cat), launch #9 of the command; kernels are serialised under ncu"
    cp $O/launches_$tag.csv profiles/launches_$tag.csv
    tail -4 profiles/ncu_walk_${tag}_summary.txt
    ;;
*) echo "unknown command $cmd"; exit 1;;
esac
