// vgen_cuda_hash160_batch_static: the hash160 batch of include/vgen_cuda.h for callers that must not allocate (code under CUDA
// graph capture, serving loops, harnesses that forbid cudaMalloc). It has no counterpart in the reference.
// It runs the SAME kernels as vgen_cuda_hash160_batch — table build, centre set-up, the walk
// kernel in dump mode, the slow pass over degenerate batches — but everything they work in (generator tables, centres, the
// batch-inversion scratch, the degenerate list, counters) lives in ONE file-scope __device__ array instead of cudaMalloc'ed
// buffers, so the call itself never calls cudaMalloc / cudaFree / cudaHostAlloc or any allocating library routine.
//
// The array is 17.45 MB and exists on every GPU that loads this library, until the process exits. One call covers at most
// VGEN_STATIC_MAX_KEYS = 2^20 keys: one walk window of 1,024 threads x 1,024 keys.
//
// Host code never uses the array's NAME as an address (in host code the name is a host-side stand-in, not device memory): once
// per device, with that device current, static_state() asks cudaGetSymbolAddress for the device address and the launches get
// pointers derived from it. Device code in this file (k_static_seed) names the array directly.
//
// Every call shares the array and finds the previous call's contents in it. The tables are read-only after set-up; centres,
// scratch and the degenerate list are fully rewritten by a call before it reads them; the counters are cleared by k_static_seed
// at the start of every call. Calls on the same device that overlap on the GPU (different streams, or different host threads on
// different streams) race on the array: the caller serialises them (one stream, or events between streams).
#include <cuda_runtime.h>
#include <mutex>
#include <stdio.h>
#include <string.h>
#include "../../include/vgen_cuda.h"
#include "host/host_util.h"
#include "kernels.h"

namespace {

enum : size_t {
    kThreads = 1024, kHalf = 512,                       // geometry of the one window: 4 CTAs of VG_BLOCK threads, 2 x 512 keys per thread
    kWindow = kThreads * 2 * kHalf,                     // = VGEN_STATIC_MAX_KEYS
    // word offsets into the array; every region starts on a 256-byte boundary
    kOffGtable = 0,                                     // pow2[256][16] + win8[32][256][16] (ec.cuh), built once per device
    kOffGtab = kOffGtable + VG_GTABLE_WORDS,            // (i+1)*G, i < 512, built once per device
    kOffCenters = kOffGtab + kHalf * 16,                // [16][kThreads] centres of the window
    kOffScratch = kOffCenters + 16 * kThreads,          // [(kHalf+1)][8][kThreads] prefix products of the batch inversion
    kOffDegen = kOffScratch + (kHalf + 1) * 8 * kThreads,   // [kThreads] threads whose batch needs the slow pass
    kOffCounters = kOffDegen + kThreads,                // [0] hit count (unused in dump mode), [1] degenerate count, [2..3] checked build
    kOffScalars = kOffCounters + 64,                    // [0..8) scalar of the first centre, [8..16) window base, [16..32) one affine point
    kWords = kOffScalars + 64
};
static_assert(kWindow == VGEN_STATIC_MAX_KEYS, "header and kernel geometry disagree");
static_assert(kOffGtab % 64 == 0 && kOffCenters % 64 == 0 && kOffScratch % 64 == 0 && kOffDegen % 64 == 0 && kOffCounters % 64 == 0 &&
              kOffScalars % 64 == 0, "regions must start on 256-byte boundaries");
static_assert(kThreads % VG_BLOCK == 0, "whole CTAs");

static __device__ __align__(256) uint32_t g_static_work[kWords];       // 4,362,624 words = 17.45 MB per GPU

struct scalar8 { uint32_t v[8]; };

// start of every call: the two scalars the set-up kernels read from device memory, and the counters back to zero
__global__ void k_static_seed(scalar8 first_centre, scalar8 window_base) {
    const uint32_t t = threadIdx.x;
    if (t < 8) { g_static_work[kOffScalars + t] = first_centre.v[t]; g_static_work[kOffScalars + 8 + t] = window_base.v[t]; }
    if (t < 64) g_static_work[kOffCounters + t] = 0;
}

struct DeviceState {
    bool ready = false;
    uint32_t *base = nullptr;          // device address of g_static_work on this device (cudaGetSymbolAddress)
    uint32_t stride_pt[16];            // kWindow * G, the next-window point the walk kernel wants in its parameters
};
const int kMaxDevices = 64;
DeviceState g_state[kMaxDevices];
std::mutex g_state_mu;

int fail(char *err, size_t cap, int code, const char *what, cudaError_t e = cudaSuccess) {
    if (err && cap) {
        if (e != cudaSuccess) snprintf(err, cap, "%s: %s", what, cudaGetErrorString(e));
        else snprintf(err, cap, "%s", what);
    }
    return code;
}
#define SB_CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(err, err_cap, VGEN_ERR_CUDA, #call, e_); } while (0)

scalar8 limbs_of(const vg::U256 &k) { scalar8 s; k.to_limbs32(s.v); return s; }

// once per device, with that device current: the array's device address, the generator tables, the stride point
int static_state(int device, cudaStream_t s, DeviceState **out, char *err, size_t err_cap) {
    std::lock_guard<std::mutex> g(g_state_mu);
    DeviceState &D = g_state[device];
    *out = &D;
    if (D.ready) return VGEN_OK;
    cudaDeviceProp prop;
    SB_CK(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) return fail(err, err_cap, VGEN_ERR_CUDA, "not an sm_100 class device; this library is built for B200 (sm_100a) only");
    void *p = nullptr;
    SB_CK(cudaGetSymbolAddress(&p, g_static_work));
    uint32_t *base = (uint32_t *)p;
    SB_CK(vg::launch_build_tables(base + kOffGtable, base + kOffGtab, (uint32_t)kHalf, s));
    k_static_seed<<<1, 64, 0, s>>>(limbs_of(vg::U256((uint64_t)kWindow)), limbs_of(vg::U256(0)));
    SB_CK(cudaGetLastError());
    SB_CK(vg::launch_points_of_scalars(base + kOffScalars + 16, 1, base + kOffScalars, 0, base + kOffGtable, s));
    SB_CK(cudaMemcpyAsync(D.stride_pt, base + kOffScalars + 16, 64, cudaMemcpyDeviceToHost, s));
    SB_CK(cudaStreamSynchronize(s));
    D.base = base;
    D.ready = true;
    return VGEN_OK;
}

}  // namespace

extern "C" int vgen_cuda_hash160_batch_static(int device, const uint8_t start_be[32], uint32_t n, uint8_t *out20_dev, void *cuda_stream,
                                              char *err, size_t err_cap) {
    if (err && err_cap) err[0] = 0;
    if (!start_be || (!out20_dev && n)) return fail(err, err_cap, VGEN_ERR_INVALID_ARG, "start_be or out20_dev is NULL");
    if (n > VGEN_STATIC_MAX_KEYS) {
        char msg[160];
        snprintf(msg, sizeof msg, "n = %u exceeds VGEN_STATIC_MAX_KEYS = %u keys per call (the static work area is sized for one window)", n, (unsigned)VGEN_STATIC_MAX_KEYS);
        return fail(err, err_cap, VGEN_ERR_INVALID_ARG, msg);
    }
    int count = 0;
    SB_CK(cudaGetDeviceCount(&count));
    if (device < 0 || device >= count || device >= kMaxDevices) return fail(err, err_cap, VGEN_ERR_INVALID_ARG, "invalid device index");
    if (n == 0) return VGEN_OK;
    SB_CK(cudaSetDevice(device));
    cudaStream_t s = (cudaStream_t)cuda_stream;
    DeviceState *D = nullptr;
    int rc = static_state(device, s, &D, err, err_cap);
    if (rc != VGEN_OK) return rc;

    // zeros for invalid scalars (0 and >= the group order), as vgen_cuda_hash160_batch: the kernels only write valid keys
    SB_CK(cudaMemsetAsync(out20_dev, 0, (size_t)n * 20, s));
    const vg::U256 start = vg::U256::from_be(start_be);
    if (!(start <= vg::secp_order_minus1())) return VGEN_OK;
    uint64_t limit = n;
    {
        vg::U256 room = vg::secp_order_minus1().sub(start).add64(1);
        if (room.fits64() && room.w[0] < limit) limit = room.w[0];
    }
    uint32_t *base = D->base;
    k_static_seed<<<1, 64, 0, s>>>(limbs_of(start.add64(kHalf)), limbs_of(start));
    SB_CK(cudaGetLastError());
    // centres (start + H + t*2H) * G, one per thread
    SB_CK(vg::launch_points_of_scalars(base + kOffCenters, (uint32_t)kThreads, base + kOffScalars, 2ull * kHalf, base + kOffGtable, s));
    vg_scan_params P;
    memset(&P, 0, sizeof P);
    P.n_threads = (uint32_t)kThreads; P.half = (uint32_t)kHalf;
    P.centers = base + kOffCenters; P.scratch = base + kOffScratch; P.gtab = base + kOffGtab;
    memcpy(P.stride_pt, D->stride_pt, 64);
    P.limit = limit;
    P.hits = nullptr; P.hit_count = base + kOffCounters; P.hit_cap = 0;
    P.degen_list = base + kOffDegen; P.degen_count = base + kOffCounters + 1; P.degen_cap = (uint32_t)kThreads;
    P.violations = base + kOffCounters + 2;
    P.dump = out20_dev; P.dump_stride = 20;
    P.m.kind = VG_MATCH_ALL;
    SB_CK(vg::launch_walk(VG_FMT_P2PKH, VG_MODE_DUMP, P, (uint32_t)(kThreads / VG_BLOCK), s));
    // batches the walk flagged (they touch scalar 0 or n, or have no valid centre): per-key scalar multiplications, the count read on the device
    SB_CK(vg::launch_slow(VG_FMT_P2PKH, VG_MODE_DUMP, P, base + kOffScalars + 8, base + kOffGtable, 0xffffffffu, s));
    return VGEN_OK;
}

#ifdef VG_CHECKED
// checked build only (tools/sanitize.sh): the device-side bounds checks of the last call on `device`: out[0] = code of the first
// failed check (scan_core.cuh), out[1] = how many failed. Synchronises the device.
extern "C" int vgen_cuda_static_violations(int device, uint32_t out[2]) {
    if (device < 0 || device >= kMaxDevices || !out) return VGEN_ERR_INVALID_ARG;
    std::lock_guard<std::mutex> g(g_state_mu);
    if (!g_state[device].ready) { out[0] = out[1] = 0; return VGEN_OK; }
    if (cudaSetDevice(device) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess ||
        cudaMemcpy(out, g_state[device].base + kOffCounters + 2, 8, cudaMemcpyDeviceToHost) != cudaSuccess) return VGEN_ERR_CUDA;
    return VGEN_OK;
}
#endif
