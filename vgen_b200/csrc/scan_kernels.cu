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
    if (blockIdx.x == 0 && threadIdx.x == 0) vg_build_pow2(pow2, gxy, pow2 + VG_GTABLE_POW2_WORDS);      // scratch: the win8 area, built afterwards
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

// One window of the P2TR fixed-base table (geometry: scan_types.h): win[v][16] = (v << off) * G for v < cnt; v = 0 is never read
// (all-zero). One thread builds VG_TAPB_GROUP consecutive digits: the first by 8-bit windowed multiplication kept in
// Jacobian form, the others by adding B = 2^off * G, and one field inversion turns the group into affine points
// (the 24-bit table has 1.8e8 entries: an inversion per entry made its build 2.4 s). Digits of a partial top window that
// would overflow 2^256 are never read; they are filled with whatever the addition chain gives.
#define VG_TAPB_GROUP 16
__global__ void k_build_tap_table(uint32_t *win, unsigned long long cnt, uint32_t off, const uint32_t *gtable) {
    const unsigned long long n = cnt;
    const unsigned long long i0 = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) * VG_TAPB_GROUP;
    if (i0 >= n) return;
    uint32_t *tab = win;
    const uint32_t v0 = (uint32_t)i0;
    const uint32_t bit = off, wi = bit >> 5, sh = bit & 31u;      // bit < 256
    uint32_t k[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    k[wi] = v0 << sh;
    if (sh != 0 && wi + 1 < 8) k[wi + 1] = v0 >> (32u - sh);
    fe Bx, By;
    for (int l = 0; l < 8; l++) { Bx.v[l] = gtable[bit * 16 + l]; By.v[l] = gtable[bit * 16 + 8 + l]; }
    // acc = k * G, Jacobian
    const uint32_t *win8 = gtable + VG_GTABLE_POW2_WORDS;
    gej_t acc; acc.inf = true;
    fe_set_zero(acc.x); fe_set_zero(acc.y); fe_set_zero(acc.z);
    for (int b = 0; b < 32; b++) {
        const uint32_t d = (k[b >> 2] >> (8 * (b & 3))) & 0xffu;
        if (!d) continue;
        const uint32_t *e = win8 + ((size_t)b * 256 + d) * 16;
        fe ex, ey;
        for (int l = 0; l < 8; l++) { ex.v[l] = e[l]; ey.v[l] = e[8 + l]; }
        gej_t nr; gej_add_affine(nr, acc, ex, ey); acc = nr;
    }
    gej_t P[VG_TAPB_GROUP];
    fe pre[VG_TAPB_GROUP];
    fe run; fe_set_one(run);
    #pragma unroll 1
    for (int g = 0; g < VG_TAPB_GROUP; g++) {
        P[g] = acc;
        if (!acc.inf) fe_mul(run, run, acc.z);
        pre[g] = run;
        gej_t nr; gej_add_affine(nr, acc, Bx, By); acc = nr;
    }
    fe inv; fe_inv(inv, run);
    #pragma unroll 1
    for (int g = VG_TAPB_GROUP - 1; g >= 0; g--) {
        fe x, y;
        if (P[g].inf) { fe_set_zero(x); fe_set_zero(y); }
        else {
            fe zi, zi2, zi3;
            if (g > 0) { fe_mul(zi, inv, pre[g - 1]); fe_mul(inv, inv, P[g].z); } else zi = inv;
            fe_sqr(zi2, zi); fe_mul(zi3, zi2, zi);
            fe_mul(x, P[g].x, zi2); fe_mul(y, P[g].y, zi3);
            fe_canon(x); fe_canon(y);
        }
        if (i0 + g < n) {
            uint4 *o = (uint4 *)(tab + (i0 + g) * 16);
            o[0] = make_uint4(x.v[0], x.v[1], x.v[2], x.v[3]); o[1] = make_uint4(x.v[4], x.v[5], x.v[6], x.v[7]);
            o[2] = make_uint4(y.v[0], y.v[1], y.v[2], y.v[3]); o[3] = make_uint4(y.v[4], y.v[5], y.v[6], y.v[7]);
        }
    }
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
cudaError_t launch_build_tap_table(uint32_t *tab, uint32_t geom, const uint32_t *pow2, cudaStream_t s) {
    if (!vg_tap_geom_valid(geom)) return cudaErrorInvalidValue;
    const uint32_t n_win = vg_tap_windows(geom);
    for (uint32_t j = 0; j < n_win; j++) {          // one launch per window (at most 64 of them)
        uint32_t off, width, off2, width2; unsigned long long first, next;
        vg_tap_window(geom, j, &off, &width, &first);
        vg_tap_window(geom, j + 1, &off2, &width2, &next);
        const unsigned long long cnt = next - first, threads = (cnt + VG_TAPB_GROUP - 1) / VG_TAPB_GROUP;
        k_build_tap_table<<<(unsigned)((threads + 127) / 128), 128, 0, s>>>(tab + first * 16, cnt, off, pow2);      // pow2: the full generator table (pow2 + win8)
        VG_CK(cudaGetLastError());
    }
    return cudaSuccess;
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
