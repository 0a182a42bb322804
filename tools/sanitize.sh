#!/bin/bash
# compute-sanitizer over every kernel of the library: memcheck, racecheck, initcheck on the short scans of
# tools/sanitize_driver.py. Run on the GPU box (gpurun); logs go to gpurun_out/sanitizer_<tag>_<tool>.log, copy them to profiles/.
set -u
tag=${1:-r02}
O=gpurun_out
mkdir -p $O
timeout 300 python tools/sanitize_driver.py > $O/sanitizer_${tag}_plain.log 2>&1 || { echo "driver fails without the sanitizer"; tail -5 $O/sanitizer_${tag}_plain.log; exit 1; }
for tool in memcheck racecheck initcheck synccheck; do
    timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_driver.py > $O/sanitizer_${tag}_$tool.log 2>&1
    echo "$tool rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $O/sanitizer_${tag}_$tool.log | tail -1) $(grep -c SANITIZE_DRIVER_OK $O/sanitizer_${tag}_$tool.log) ok-lines"
done
