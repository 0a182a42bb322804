// Launchers implemented in scan_kernels.cu, called by the host engine.
#pragma once
#include <cuda_runtime.h>
#include "scan_types.h"

#ifndef VG_BLOCK
#define VG_BLOCK 256
#endif
// resident CTAs per SM the walk kernel is compiled for (__launch_bounds__) and the full-size window is sized for
#ifndef VG_CTAS_PER_SM
#define VG_CTAS_PER_SM 2
#endif

// words of the generator table buffer (pow2[256][16] + win8[32][256][16]); must match ec.cuh
#ifndef VG_GTABLE_WORDS
#define VG_GTABLE_WORDS (256 * 16 + 32 * 256 * 16)
#endif
namespace vg {
cudaError_t launch_build_tables(uint32_t *pow2, uint32_t *gtab, uint32_t half, cudaStream_t s);
cudaError_t launch_points_of_scalars(uint32_t *out, uint32_t n, const uint32_t *base8_dev, unsigned long long step,
                                     const uint32_t *pow2, cudaStream_t s);
// P2TR: fixed-base table tab[256/bits][1 << bits][16] (v * 2^(bits*j) * G) and the stage-2 launch (taproot.cuh)
cudaError_t launch_build_tap_table(uint32_t *tab, uint32_t bits, const uint32_t *pow2, cudaStream_t s);
cudaError_t launch_taptweak(int mode, const vg_scan_params &P, cudaStream_t s);
cudaError_t launch_dump_points_x(const vg_scan_params &P, cudaStream_t s);      // untweaked internal keys (x-only) of the window
cudaError_t launch_walk(int fmt, int mode, const vg_scan_params &P, uint32_t blocks, cudaStream_t s);
cudaError_t launch_slow(int fmt, int mode, const vg_scan_params &P, const uint32_t *base8_dev, const uint32_t *pow2,
                        uint32_t n_batches, cudaStream_t s);
cudaError_t launch_walk_fmt0(int mode, const vg_scan_params &P, uint32_t blocks, cudaStream_t s);
cudaError_t launch_slow_fmt0(int mode, const vg_scan_params &P, const uint32_t *base8_dev, const uint32_t *pow2, uint32_t n_batches, cudaStream_t s);
cudaError_t launch_walk_fmt1(int mode, const vg_scan_params &P, uint32_t blocks, cudaStream_t s);
cudaError_t launch_slow_fmt1(int mode, const vg_scan_params &P, const uint32_t *base8_dev, const uint32_t *pow2, uint32_t n_batches, cudaStream_t s);
cudaError_t launch_walk_fmt2(int mode, const vg_scan_params &P, uint32_t blocks, cudaStream_t s);
cudaError_t launch_slow_fmt2(int mode, const vg_scan_params &P, const uint32_t *base8_dev, const uint32_t *pow2, uint32_t n_batches, cudaStream_t s);
cudaError_t launch_walk_fmt3(int mode, const vg_scan_params &P, uint32_t blocks, cudaStream_t s);
cudaError_t launch_slow_fmt3(int mode, const vg_scan_params &P, const uint32_t *base8_dev, const uint32_t *pow2, uint32_t n_batches, cudaStream_t s);
cudaError_t launch_walk_fmt4(int mode, const vg_scan_params &P, uint32_t blocks, cudaStream_t s);
cudaError_t launch_slow_fmt4(int mode, const vg_scan_params &P, const uint32_t *base8_dev, const uint32_t *pow2, uint32_t n_batches, cudaStream_t s);
cudaError_t launch_walk_fmt5(int mode, const vg_scan_params &P, uint32_t blocks, cudaStream_t s);
cudaError_t launch_slow_fmt5(int mode, const vg_scan_params &P, const uint32_t *base8_dev, const uint32_t *pow2, uint32_t n_batches, cudaStream_t s);
}
