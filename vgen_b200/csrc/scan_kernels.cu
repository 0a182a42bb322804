// __global__ wrappers and launchers for the B200 (sm_100a) scan path. See scan_core.cuh for the
// per-thread algorithm and DESIGN.md for the data layout. Supersedes the reference's WGSL entry points
// init_table (src/shaders/init.wgsl:3-10), main / compute_jacobian / batch_normalize_hash
// (src/shaders/search.wgsl:2-152) and their dispatch code (src/gpu.rs:502-518,535-600).
#include <cuda_runtime.h>
#include "scan_core.cuh"
#include "taproot.cuh"
#include "kernels.h"

namespace vg {

__global__ void k_build_pow2(uint32_t *pow2) {
    const uint32_t gxy[16] = {
        0x16F81798u, 0x59F2815Bu, 0x2DCE28D9u, 0x029BFCDBu, 0xCE870B07u, 0x55A06295u, 0xF9DCBBACu, 0x79BE667Eu,
        0xFB10D4B8u, 0x9C47D08Fu, 0xA6855419u, 0xFD17B448u, 0x0E1108A8u, 0x5DA4FBFCu, 0x26A3C465u, 0x483ADA77u};
    if (blockIdx.x == 0 && threadIdx.x == 0) vg_build_pow2(pow2, gxy);
}

// out[16][n] (limb-major) column t = (base + t*step) * G; all-zero column when the scalar is invalid
__global__ void k_points_of_scalars(uint32_t *out, uint32_t n, const uint32_t *base8, unsigned long long step, const uint32_t *pow2) {
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    uint32_t b[8], k[8];
    for (int i = 0; i < 8; i++) b[i] = base8[i];
    // k = base + t*step (t*step < 2^64 by construction)
    scalar_add_u64(k, b, (unsigned long long)t * step);
    vg_point_of_scalar(out, n, t, k, pow2);
}
// win8[j][v][16] = (v << 8j) * G behind the pow2 part of the generator table (ec.cuh: scalar_mul_G)
__global__ void k_build_win8(uint32_t *gtable) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= 32 * 256) return;
    uint32_t j = i >> 8, v = i & 255u;
    uint32_t k[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    k[j >> 2] = v << (8 * (j & 3));
    fe x, y;
    bool ok = v != 0 && scalar_mul_pow2(x, y, k, gtable);
    if (!ok) { fe_set_zero(x); fe_set_zero(y); }
    uint32_t *o = gtable + VG_GTABLE_POW2_WORDS + (size_t)i * 16;
    for (int l = 0; l < 8; l++) { o[l] = x.v[l]; o[8 + l] = y.v[l]; }
}
// gtab[i][16] = (i+1)*G, point-major (x limbs then y limbs)
__global__ void k_build_gtab(uint32_t *gtab, uint32_t half, const uint32_t *pow2) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= half) return;
    uint32_t k[8] = {i + 1, 0, 0, 0, 0, 0, 0, 0};
    fe x, y;
    scalar_mul_pow2(x, y, k, pow2);
    for (int l = 0; l < 8; l++) { gtab[i * 16 + l] = x.v[l]; gtab[i * 16 + 8 + l] = y.v[l]; }
}

// tab[j][v][16] = (v << (bits*j)) * G for the P2TR tweak addition; v = 0 entries are never read (all-zero)
__global__ void k_build_tap_table(uint32_t *tab, uint32_t bits, const uint32_t *pow2) {
    const unsigned long long n = (unsigned long long)vg_tap_windows(bits) << bits;
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t j = (uint32_t)(i >> bits), v = (uint32_t)(i & ((1ull << bits) - 1));
    uint32_t k[8];
    fe x, y;
    bool ok = vg_tap_digit_scalar(k, j, v, bits) && scalar_mul_pow2(x, y, k, pow2);
    if (!ok) { fe_set_zero(x); fe_set_zero(y); }
    uint4 *o = (uint4 *)(tab + i * 16);
    o[0] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]); o[1] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]);
    o[2] = make_uint4(y.v[0], y.v[1], y.v[2], y.v[3]); o[3] = make_uint4(y.v[4], y.v[5], y.v[6], y.v[7]);
}

// P2TR internal keys as the reference's dispatch_p2tr / await_result_p2tr return them (src/gpu.rs:667-807): the
// big-endian x of the even-Y point, before the TapTweak. Reads stage 1's point buffer.
__global__ void k_dump_points_x(const uint32_t *points, unsigned long long limit, uint8_t *out, uint32_t stride) {
    unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= limit) return;
    for (uint32_t b = 0; b < stride && b < 32; b++) {
        uint32_t w = points[i * 16 + (7 - (b >> 2))];
        out[i * stride + b] = (uint8_t)(w >> (8 * (3 - (b & 3))));
    }
}

#define VG_CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

cudaError_t launch_build_tables(uint32_t *pow2, uint32_t *gtab, uint32_t half, cudaStream_t s) {
    k_build_pow2<<<1, 32, 0, s>>>(pow2);
    VG_CK(cudaGetLastError());
    k_build_win8<<<64, 128, 0, s>>>(pow2);
    VG_CK(cudaGetLastError());
    k_build_gtab<<<(half + 127) / 128, 128, 0, s>>>(gtab, half, pow2);
    return cudaGetLastError();
}
cudaError_t launch_build_tap_table(uint32_t *tab, uint32_t bits, const uint32_t *pow2, cudaStream_t s) {
    const unsigned long long n = (unsigned long long)((256u + bits - 1u) / bits) << bits;
    k_build_tap_table<<<(unsigned)((n + 127) / 128), 128, 0, s>>>(tab, bits, pow2);
    return cudaGetLastError();
}
cudaError_t launch_dump_points_x(const vg_scan_params &P, cudaStream_t s) {
    if (P.limit == 0) return cudaSuccess;
    k_dump_points_x<<<(unsigned)((P.limit + 255) / 256), 256, 0, s>>>(P.points, P.limit, P.dump, P.dump_stride);
    return cudaGetLastError();
}
cudaError_t launch_points_of_scalars(uint32_t *out, uint32_t n, const uint32_t *base8_dev, unsigned long long step,
                                     const uint32_t *pow2, cudaStream_t s) {
    k_points_of_scalars<<<(n + 127) / 128, 128, 0, s>>>(out, n, base8_dev, step, pow2);
    return cudaGetLastError();
}

cudaError_t launch_walk(int fmt, int mode, const vg_scan_params &P, uint32_t blocks, cudaStream_t s) {
    switch (fmt) {
        case 0: return launch_walk_fmt0(mode, P, blocks, s); case 1: return launch_walk_fmt1(mode, P, blocks, s);
        case 2: return launch_walk_fmt2(mode, P, blocks, s); case 3: return launch_walk_fmt3(mode, P, blocks, s);
        case 4: return launch_walk_fmt4(mode, P, blocks, s); case 5: return launch_walk_fmt5(mode, P, blocks, s);
    }
    return cudaErrorInvalidValue;
}
cudaError_t launch_slow(int fmt, int mode, const vg_scan_params &P, const uint32_t *base8_dev, const uint32_t *pow2,
                        uint32_t n_batches, cudaStream_t s) {
    switch (fmt) {
        case 0: return launch_slow_fmt0(mode, P, base8_dev, pow2, n_batches, s); case 1: return launch_slow_fmt1(mode, P, base8_dev, pow2, n_batches, s);
        case 2: return launch_slow_fmt2(mode, P, base8_dev, pow2, n_batches, s); case 3: return launch_slow_fmt3(mode, P, base8_dev, pow2, n_batches, s);
        case 4: return launch_slow_fmt4(mode, P, base8_dev, pow2, n_batches, s); case 5: return launch_slow_fmt5(mode, P, base8_dev, pow2, n_batches, s);
    }
    return cudaErrorInvalidValue;
}

}  // namespace vg
