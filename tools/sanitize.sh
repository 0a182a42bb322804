#!/bin/bash
# Memory-safety evidence for every kernel of the library.
#
# compute-sanitizer is CLOSED on this GPU pool ("runs under it have left GPUs needing a reset": gpurun_out/sanitizer_r02_memcheck.log,
# kept as profiles/sanitizer_r02_closed.log), so the checks are the library's own:
#   * libvgen_cuda_checked.so = the same sources built with -DVG_CHECKED: every index the kernels compute (thread id, scratch / table /
#     hit / degenerate-list / point-buffer / DFA / fixed-base-table index, dynamic shared-memory size) is asserted on the device; a
#     failed assert makes the scan return an error with the check's code (scan_core.cuh VG_ASSERT, engine.cpp harvest);
#   * VGEN_GUARD=1 puts a 256-byte canary behind every device buffer; vgen_cuda_stats reports how many were damaged (out-of-bounds WRITES);
#   * tools/sanitize_driver.py drives all kernels through all matcher modes, incl. k_slow (ranges with scalar 0 and n-1) and
#     k_taptweak with a range end inside a warp and inside a CTA; the results of the same scans are compared with the oracle by the GPU
#     test suite.
# Build the checked library first (here or on the box): make -C vgen_b200/csrc -j8 BUILD=build_checked OUT=../libvgen_cuda_checked.so EXTRA=-DVG_CHECKED ../libvgen_cuda_checked.so
set -u
tag=${1:-r02}
O=gpurun_out
mkdir -p $O
[ -f vgen_b200/libvgen_cuda_checked.so ] || make -C vgen_b200/csrc -j8 BUILD=build_checked OUT=../libvgen_cuda_checked.so EXTRA=-DVG_CHECKED ../libvgen_cuda_checked.so > $O/checked_build_$tag.log 2>&1
VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda_checked.so VGEN_GUARD=1 timeout 900 python tools/sanitize_driver.py > $O/checked_${tag}.log 2>&1
echo "checked build rc=$? $(grep -c SANITIZE_DRIVER_OK $O/checked_${tag}.log) ok-lines; $(grep '^stats' $O/checked_${tag}.log | tail -1)"
# the default geometry (296 CTAs, full-size windows) through the checked library as well: one window per format
VGEN_CUDA_LIB=$PWD/vgen_b200/libvgen_cuda_checked.so VGEN_GUARD=1 timeout 900 python tools/sanitize_driver.py --full > $O/checked_full_${tag}.log 2>&1
echo "checked build, full-size windows rc=$? $(grep -c SANITIZE_DRIVER_OK $O/checked_full_${tag}.log) ok-lines; $(grep '^stats' $O/checked_full_${tag}.log | tail -1)"
