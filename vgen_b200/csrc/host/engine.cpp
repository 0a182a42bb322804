// Host driver of the B200 scan path (see engine.h).
//
// Window schedule. With G global lanes (shards x GPUs x lanes per GPU) the keyspace [start, start+N) is cut into a fixed
// sequence of windows: one small window per lane at each ramp level (so short ranges and dense patterns finish without
// computing a full-size window), then full-size windows handed out round-robin: lane g takes windows g, g+G, g+2G, ...
// Each lane therefore walks with a disjoint start scalar and a common stride of G*W keys (the walk kernel adds the stride
// point (G*W)*G to every centre), there is no device-to-device traffic, and hits come back in ascending key order.
//
// Threads. Every device of the context is driven by its own host thread (device_worker): it issues the windows of the
// device's lanes, waits for them, copies the hit records out through the slot's pinned staging buffer and hands them to a
// pool of confirm workers (re-encode + full DFA + text). The calling thread only merges finished windows in sequence order,
// calls the progress callback and polls the stop flag. A device whose CUDA call fails (or VGEN_FAULT_INJECT=<dev>:<n>) drops
// out: its unfinished windows go back into the queue and its lanes' future windows are taken by the surviving devices, each
// picking the lowest outstanding window first (the price is a centre set-up per adopted window). The reference instead falls
// back to the CPU scanner (src/lib.rs:727-746,1185-1197), which BASELINE.json's north_star rules out.
#include "engine.h"
#include <algorithm>
#include <chrono>
#include <map>
#include <string.h>
#include <stdio.h>
#include <stdlib.h>

namespace vg {

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { err = std::string(#call) + ": " + cudaGetErrorString(e_); return false; } } while (0)

static const int kCounterWords = 4;
static const uint32_t kEagerTextHits = 1u << 16;   // windows with at most this many device hits get their text records built by the confirm job
static const uint32_t kStageRecords = 1u << 20;    // pinned staging records per slot (40 MB)
static const uint64_t kLookaheadPerLane = 8;       // a lane does not start a window more than this many rounds ahead of the merge

// ---- worker pool ----
WorkerPool::WorkerPool(unsigned n) {
    for (unsigned i = 0; i < n; i++) threads_.emplace_back([this] { run(); });
}
WorkerPool::~WorkerPool() {
    { std::lock_guard<std::mutex> g(mu_); stop_ = true; }
    cv_.notify_all();
    for (auto &t : threads_) t.join();
}
void WorkerPool::run() {
    for (;;) {
        std::packaged_task<HitBatch()> job;
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [this] { return stop_ || !q_.empty(); });
            if (q_.empty()) return;
            job = std::move(q_.front()); q_.pop_front();
        }
        job();
    }
}
std::future<HitBatch> WorkerPool::submit(std::function<HitBatch()> fn) {
    std::packaged_task<HitBatch()> task(std::move(fn));
    std::future<HitBatch> f = task.get_future();
    { std::lock_guard<std::mutex> g(mu_); q_.push_back(std::move(task)); }
    cv_.notify_one();
    return f;
}

Engine::~Engine() { destroy(); }

static const size_t kGuardBytes = 256;
cudaError_t Engine::dmalloc(void **p, size_t bytes, int device) {
    if (!guard_) return cudaMalloc(p, bytes);
    const size_t padded = (bytes + 15) & ~(size_t)15;
    cudaError_t e = cudaMalloc(p, padded + kGuardBytes);
    if (e != cudaSuccess) return e;
    e = cudaMemset((uint8_t *)*p + padded, 0xA5, kGuardBytes);
    std::lock_guard<std::mutex> g(guards_mu_);
    guards_.push_back({device, (uint8_t *)*p + padded});
    return e;
}
std::pair<uint64_t, uint64_t> Engine::check_guards() {
    std::lock_guard<std::mutex> g(guards_mu_);
    uint64_t bad = 0;
    uint8_t host[kGuardBytes];
    for (auto &G : guards_) {
        if (cudaSetDevice(G.device) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess ||
            cudaMemcpy(host, G.at, kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess) { bad++; continue; }
        for (size_t i = 0; i < kGuardBytes; i++) if (host[i] != 0xA5) { bad++; break; }
    }
    return {(uint64_t)guards_.size(), bad};
}

int Engine::lanes_per_device() {
    const char *e = getenv("VGEN_LANES_PER_GPU");       // tuning override
    int n = e ? atoi(e) : 2;
    return n < 1 ? 1 : (n > 4 ? 4 : n);
}

void Engine::destroy() {
    pool_.reset();
    for (auto &L : lanes_) {
        cudaSetDevice(L.device);
        if (L.stream) cudaStreamSynchronize(L.stream);
        for (auto &S : L.slot) {
            if (S.d_hits) cudaFree(S.d_hits);
            if (S.d_counters) cudaFree(S.d_counters);
            if (S.d_degen) cudaFree(S.d_degen);
            if (S.d_base8) cudaFree(S.d_base8);
            if (S.h_counters) cudaFreeHost(S.h_counters);
            if (S.h_hits) cudaFreeHost(S.h_hits);
            if (S.done) cudaEventDestroy(S.done);
        }
        for (void *p : {(void *)L.d_centers, (void *)L.d_scratch, (void *)L.d_base8, (void *)L.d_stride_pts, (void *)L.d_points})
            if (p) cudaFree(p);
        if (L.own_stream && L.stream) cudaStreamDestroy(L.stream);
        if (L.copy_stream) cudaStreamDestroy(L.copy_stream);
    }
    lanes_.clear();
    for (auto &D : devices_) {
        cudaSetDevice(D.device);
        for (void *p : {(void *)D.d_pow2, (void *)D.d_gtab, (void *)D.d_tap_table, (void *)D.d_dfa_table, (void *)D.d_dfa_accept})
            if (p) cudaFree(p);
    }
    devices_.clear();
}

bool Engine::create(const int *devices, int n_devices, uint32_t hint, std::string &err) {
    int count = 0;
    cudaError_t ce = cudaGetDeviceCount(&count);
    if (ce != cudaSuccess || count == 0) { err = std::string("no CUDA device: ") + cudaGetErrorString(ce); return false; }
    std::vector<int> devs;
    if (devices && n_devices > 0) devs.assign(devices, devices + n_devices);
    else devs.push_back(0);
    for (int d : devs) if (d < 0 || d >= count) { err = "invalid device index " + std::to_string(d); return false; }
    half_ = 512;
    guard_ = getenv("VGEN_GUARD") != nullptr;
    int min_sm = 1 << 30;
    for (int d : devs) {
        cudaDeviceProp prop;
        CK(cudaGetDeviceProperties(&prop, d));
        if (prop.major < 10) { err = std::string("device ") + prop.name + " is not sm_100 class; this library is built for B200 (sm_100a) only"; return false; }
        Device D; D.device = d; D.sm_count = prop.multiProcessorCount; D.name = prop.name;
        min_sm = std::min(min_sm, D.sm_count);
        // Several lanes (stream + centres + scratch + hit buffers) per GPU: a window's warps do not finish together
        // (ncu r01: 12.9 of 16 resident warps active on average), and a second stream's next window fills the CTA slots the
        // early finishers free. Measured on B200: 1 lane 5,819, 2 lanes 6,094, 3 lanes 6,085 Mkeys/s (profiles/overlap_r01.log).
        for (int k = 0; k < lanes_per_device(); k++) {
            Lane L; L.dev = (int)devices_.size(); L.device = d;
            D.lanes.push_back((int)lanes_.size());
            lanes_.push_back(L);
        }
        devices_.push_back(D);
    }
    levels_ = compute_levels(min_sm, hint, half_, VG_BLOCK, VG_CTAS_PER_SM);
    const Level &F = levels_.back();
    // Hit buffers hold at most 2^24 records per slot (671 MB), not one per key of a full window (77.6 M x 40 B = 3.1 GB per
    // slot, 12.4 GB per GPU in round 1). A window that reports more (a dense pattern on a full-size window: host-bound anyway)
    // is re-run in pieces of the largest ramp level that fits (harvest). VGEN_HIT_CAP shrinks the buffers (tests).
    uint64_t cap = std::min<uint64_t>(F.window, 1ull << 24);
    if (const char *e = getenv("VGEN_HIT_CAP")) { uint64_t v = strtoull(e, nullptr, 10); if (v) cap = std::min<uint64_t>(cap, v); }
    cap = std::max<uint64_t>(cap, levels_[0].window);
    hit_cap_ = (uint32_t)cap;
    rescan_level_ = 0;
    for (int lv = 0; lv < (int)levels_.size(); lv++) if (levels_[lv].window <= cap) rescan_level_ = lv;
    for (auto &D : devices_) {
        CK(cudaSetDevice(D.device));
        CK(dmalloc((void **)&D.d_pow2, (size_t)VG_GTABLE_WORDS * 4, D.device));       // pow2[256] + win8[32][256] (ec.cuh)
        CK(dmalloc((void **)&D.d_gtab, (size_t)half_ * 16 * 4, D.device));
    }
    for (auto &L : lanes_) {
        CK(cudaSetDevice(L.device));
        CK(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&L.copy_stream, cudaStreamNonBlocking));
        CK(dmalloc((void **)&L.d_centers, (size_t)16 * F.threads * 4, L.device));
        CK(dmalloc((void **)&L.d_scratch, (size_t)(half_ + 1) * 8 * F.threads * 4, L.device));
        CK(dmalloc((void **)&L.d_base8, 64, L.device));
        CK(dmalloc((void **)&L.d_stride_pts, 16 * 4, L.device));
        for (auto &S : L.slot) {
            CK(dmalloc((void **)&S.d_hits, (size_t)hit_cap_ * sizeof(vg_hit_rec), L.device));
            CK(dmalloc((void **)&S.d_counters, kCounterWords * 4, L.device));
            CK(dmalloc((void **)&S.d_degen, (size_t)F.threads * 4, L.device));
            CK(dmalloc((void **)&S.d_base8, 32, L.device));
            CK(cudaHostAlloc(&S.h_counters, kCounterWords * 4, cudaHostAllocDefault));
            CK(cudaHostAlloc(&S.h_hits, (size_t)std::min(hit_cap_, kStageRecords) * sizeof(vg_hit_rec), cudaHostAllocDefault));
            CK(cudaEventCreateWithFlags(&S.done, cudaEventDisableTiming));
        }
    }
    // generator tables: once per device (its lanes share them), all devices building at the same time
    for (auto &D : devices_) {
        CK(cudaSetDevice(D.device));
        CK(launch_build_tables(D.d_pow2, D.d_gtab, half_, lanes_[D.lanes[0]].stream));
        launches += 3;
    }
    for (auto &D : devices_) { CK(cudaSetDevice(D.device)); CK(cudaStreamSynchronize(lanes_[D.lanes[0]].stream)); }
    // stride of a lane walking alone (digest_batch): W*G per level
    {
        Lane &L = lanes_[0];
        CK(cudaSetDevice(L.device));
        stride1_.assign(levels_.size(), {});
        for (size_t lv = 0; lv < levels_.size(); lv++) {
            uint32_t k8[8]; U256(levels_[lv].window).to_limbs32(k8);
            CK(cudaMemcpyAsync(L.d_base8, k8, 32, cudaMemcpyHostToDevice, L.stream));
            CK(launch_points_of_scalars(L.d_stride_pts, 1, L.d_base8, 0, devices_[L.dev].d_pow2, L.stream));
            launches++;
            CK(cudaMemcpyAsync(stride1_[lv].data(), L.d_stride_pts, 64, cudaMemcpyDeviceToHost, L.stream));
            CK(cudaStreamSynchronize(L.stream));
        }
    }
    unsigned hw = std::thread::hardware_concurrency();
    pool_.reset(new WorkerPool(std::max(2u, std::min(hw ? hw : 4u, 32u))));
    return true;
}

bool Engine::set_stream(int lane, cudaStream_t s, std::string &err) {
    if (lane < 0 || lane >= (int)lanes_.size()) { err = "invalid lane"; return false; }
    Lane &L = lanes_[lane];
    CK(cudaSetDevice(L.device));
    CK(cudaStreamSynchronize(L.stream));
    if (L.own_stream) cudaStreamDestroy(L.stream);
    L.stream = s; L.own_stream = false;
    return true;
}

bool Engine::set_shard(uint32_t index, uint32_t count, std::string &err) {
    if (count == 0 || index >= count) { err = "invalid shard index/count"; return false; }
    shard_index_ = index; shard_count_ = count; stride_lanes_ = 0;
    return true;
}

bool Engine::upload_dfa(std::string &err) {
    const Dfa &d = lp_.dfa;
    // the kernel stages both arrays word-wise: whole words, tail cleared
    const size_t tb = (d.table.size() * 2 + 7) & ~(size_t)3, ab = (d.accept_end.size() + 7) & ~(size_t)3;
    std::vector<uint8_t> tbuf(tb, 0), abuf(ab, 0);
    memcpy(tbuf.data(), d.table.data(), d.table.size() * 2);
    memcpy(abuf.data(), d.accept_end.data(), d.accept_end.size());
    for (auto &D : devices_) {
        if (D.failed) continue;
        CK(cudaSetDevice(D.device));
        // buffers are kept and reused when the new tables fit: cudaFree / cudaMalloc synchronise the device
        if (D.dfa_table_cap < tb) { if (D.d_dfa_table) cudaFree(D.d_dfa_table); D.d_dfa_table = nullptr; CK(dmalloc((void **)&D.d_dfa_table, std::max<size_t>(tb, 4096), D.device)); D.dfa_table_cap = std::max<size_t>(tb, 4096); }
        if (D.dfa_accept_cap < ab) { if (D.d_dfa_accept) cudaFree(D.d_dfa_accept); D.d_dfa_accept = nullptr; CK(dmalloc((void **)&D.d_dfa_accept, std::max<size_t>(ab, 1024), D.device)); D.dfa_accept_cap = std::max<size_t>(ab, 1024); }
        CK(cudaMemcpy(D.d_dfa_table, tbuf.data(), tb, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(D.d_dfa_accept, abuf.data(), ab, cudaMemcpyHostToDevice));
    }
    dfa_uploaded_ = true;
    return true;
}

// Geometry of the P2TR fixed-base table (scan_types.h), chosen per job. Measured on B200 with the round-2 stage
// (profiles/tap_geometries_r02.log, tap_signed_r02.log; round 1: tap_bits_r01.log), P2TR ^bc1pcat, Mkeys/s:
//   16 bits            16 windows x 65,536 entries          67 MB   built in 6 ms    564
//   s17.1 (signed)     15 windows, 1 x 18 + 14 x 17 bits    67 MB            5 ms    587      <- jobs below 2^30 keys
//   24 bits            11 windows x 2^24                   11.8 GB         0.15 s    787
//   s25.6 (signed)     10 windows, 6 x 26 + 4 x 25 bits    17.2 GB          0.3 s    840      <- jobs of 2^30 .. 2^36 keys
//   26 bits            10 windows x 2^26                   42.9 GB         0.45 s    856      (fallback when 112 GB are not free)
//   s28.4 (signed)      9 windows, 4 x 29 + 5 x 28 bits   111.7 GB          1.2 s    905      <- jobs of at least 2^36 keys
// The entries are gathered at random from DRAM either way (ncu: even the 64 MiB table is flushed out of L2 by the window's point
// stream), so a bigger table only saves additions; signed digits halve the entries per window at the price of a conditional
// negation (2 % at equal window count). VGEN_TAP_BITS forces a geometry: 4..26, or sW.H for the signed one with H windows of
// W + 1 bits and the rest of W bits.
static const uint32_t kTapShort = VG_TAP_GEOM_SIGNED(17, 1), kTapMid = VG_TAP_GEOM_SIGNED(25, 6), kTapLong = VG_TAP_GEOM_SIGNED(28, 4);
static uint32_t tap_bits_for(uint64_t keys_hint) {
    if (const char *e = getenv("VGEN_TAP_BITS")) {
        if (e[0] == 's' || e[0] == 'S') {
            unsigned W = 0, H = 0;
            if (sscanf(e + 1, "%u.%u", &W, &H) >= 1 && W < 256 && H < 256 && vg_tap_geom_valid(VG_TAP_GEOM_SIGNED(W, H))) return VG_TAP_GEOM_SIGNED(W, H);
            return kTapShort;
        }
        long v = strtol(e, nullptr, 10);
        return (uint32_t)(v < 4 ? 4 : v > 26 ? 26 : v);
    }
    return keys_hint >= (1ull << 36) ? kTapLong : keys_hint >= (1ull << 30) ? kTapMid : kTapShort;
}
// a table's size is also its rank: of two geometries the bigger one is the faster one (table above)
static size_t tap_table_bytes(uint32_t geom) { return geom ? (size_t)vg_tap_entries(geom) * 64 : 0; }

// P2TR needs two more buffers: the stage-1 -> stage-2 point buffer of every lane (64 B per window key) and the fixed-base
// table of the tweak stage (one per GPU, shared by its lanes). Allocated and built on first use; a bigger table replaces
// a smaller one when a big job comes, never the other way round.
bool Engine::ensure_taproot(uint64_t keys_hint, std::string &err) {
    const Level &F = levels_.back();
    uint32_t want = tap_bits_for(keys_hint);
    for (auto &L : lanes_) {
        if (L.d_points || devices_[L.dev].failed) continue;
        CK(cudaSetDevice(L.device));
        CK(dmalloc((void **)&L.d_points, (size_t)F.window * 64, L.device));
    }
    if (tap_bits_ != 0 && (tap_table_bytes(tap_bits_) >= tap_table_bytes(want) || getenv("VGEN_TAP_BITS"))) return true;
    // a big table needs room: on a crowded device step down to the next smaller one of the list
    const uint32_t steps[] = {want, 26u, kTapMid, kTapShort};
    for (auto &D : devices_) {
        if (D.failed) continue;
        size_t free_b = 0, total_b = 0;
        CK(cudaSetDevice(D.device));
        CK(cudaMemGetInfo(&free_b, &total_b));
        const size_t held = tap_table_bytes(tap_bits_);                       // freed before the new one is allocated
        for (uint32_t g : steps) {
            if (tap_table_bytes(g) > tap_table_bytes(want)) continue;
            want = g;
            const size_t spare = tap_table_bytes(g) > (1ull << 30) ? (8ull << 30) : (1ull << 30);
            if (free_b + held >= tap_table_bytes(g) + spare) break;
        }
    }
    if (tap_bits_ != 0 && tap_table_bytes(tap_bits_) >= tap_table_bytes(want)) return true;
    const auto t0 = std::chrono::steady_clock::now();
    for (auto &L : lanes_) { if (devices_[L.dev].failed) continue; CK(cudaSetDevice(L.device)); CK(cudaStreamSynchronize(L.stream)); }
    for (auto &D : devices_) {
        if (D.failed) continue;
        CK(cudaSetDevice(D.device));
        if (D.d_tap_table) { CK(cudaFree(D.d_tap_table)); D.d_tap_table = nullptr; }
        CK(dmalloc((void **)&D.d_tap_table, tap_table_bytes(want), D.device));
        CK(launch_build_tap_table(D.d_tap_table, want, D.d_pow2, lanes_[D.lanes[0]].stream));
        launches++;
    }
    for (auto &D : devices_) { if (D.failed) continue; CK(cudaSetDevice(D.device)); CK(cudaStreamSynchronize(lanes_[D.lanes[0]].stream)); }
    tap_bits_ = want;
    if (getenv("VGEN_TRACE"))
        fprintf(stderr, "[vgen trace] tap table 0x%x (%u windows, %.1f GB) %9.3f ms\n", want, vg_tap_windows(want), tap_table_bytes(want) / 1e9, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    return true;
}

bool Engine::set_pattern(int format, const std::string &pattern, bool ci, std::string &err) {
    // the same pattern again (scan after scan through scan_gpu_with_runner): nothing to lower, nothing to upload
    if (has_pattern_ && dfa_uploaded_ && lp_.format == format && lp_.pattern == pattern && lp_.case_insensitive == ci) return true;
    dfa_uploaded_ = false;
    LoweredPattern lp;
    if (!lower_pattern(format, pattern, ci, lp, err)) return false;
    lp_ = lp;
    has_pattern_ = true;
    return upload_dfa(err);
}

// stride points of every level for the current global lane count: (G * W_level) * G. The same for every lane: computed once
// on the first healthy device.
bool Engine::compute_stride_points(std::string &err) {
    uint32_t G = shard_count_ * (uint32_t)lanes_.size();
    if (stride_lanes_ == G) return true;
    Lane *L0 = nullptr;
    for (auto &L : lanes_) if (!devices_[L.dev].failed) { L0 = &L; break; }
    if (!L0) { err = "every device of this context has failed"; return false; }
    CK(cudaSetDevice(L0->device));
    std::vector<std::array<uint32_t, 16>> pts(levels_.size());
    for (size_t lv = 0; lv < levels_.size(); lv++) {
        // every level uses its own G*W stride; only the final level's next centre is ever consumed
        U256 k = U256(levels_[lv].window).mul64(G);
        uint32_t k8[8]; k.to_limbs32(k8);
        CK(cudaMemcpyAsync(L0->d_base8, k8, 32, cudaMemcpyHostToDevice, L0->stream));
        CK(launch_points_of_scalars(L0->d_stride_pts, 1, L0->d_base8, 0, devices_[L0->dev].d_pow2, L0->stream));
        launches++;
        CK(cudaMemcpyAsync(pts[lv].data(), L0->d_stride_pts, 64, cudaMemcpyDeviceToHost, L0->stream));
        CK(cudaStreamSynchronize(L0->stream));
    }
    for (auto &L : lanes_) { L.stride_pt = pts; L.centers_level = -1; }
    stride_lanes_ = G;
    return true;
}

bool Engine::setup_centers(Lane &L, int level, const U256 &window_base, std::string &err) {
    const Level &V = levels_[level];
    U256 first = window_base.add64(half_);
    uint32_t k8[8]; first.to_limbs32(k8);
    CK(cudaMemcpyAsync(L.d_base8, k8, 32, cudaMemcpyHostToDevice, L.stream));
    CK(launch_points_of_scalars(L.d_centers, V.threads, L.d_base8, 2ull * half_, devices_[L.dev].d_pow2, L.stream));
    launches++; h2d_bytes += 32;
    L.centers_level = level; L.centers_base = window_base;
    return true;
}

void Engine::fill_params(Lane &L, Slot &S, const uint32_t *stride16, vg_scan_params &P) {
    const Level &V = levels_[S.task.level];
    const Device &D = devices_[L.dev];
    memset(&P, 0, sizeof P);
    P.n_threads = V.threads; P.half = half_;
    P.centers = L.d_centers; P.scratch = L.d_scratch; P.gtab = D.d_gtab;
    memcpy(P.stride_pt, stride16, 64);
    P.limit = S.task.limit;
    P.hits = S.d_hits; P.hit_count = S.d_counters; P.hit_cap = hit_cap_;
    P.degen_list = S.d_degen; P.degen_count = S.d_counters + 1; P.degen_cap = V.threads;
    P.points = L.d_points; P.tap_table = D.d_tap_table; P.tap_bits = tap_bits_;
    S.task.base.to_limbs32(P.tap_base);
    P.violations = S.d_counters + 2;
    P.m = lp_.dev;
    P.m.dfa.table = D.d_dfa_table; P.m.dfa.accept = D.d_dfa_accept;
    P.dfa_smem_bytes = (P.m.kind == VG_MATCH_DFA || P.m.confirm_dfa) ? dfa_smem_bytes(P.m.dfa) : 0;
}

// launches the window described by S.task on lane L (the centres must stand at S.task.base); the counters come back through
// S.h_counters, completion through S.done
bool Engine::launch_window(Lane &L, Slot &S, int fmt, int mode, uint8_t *dump, uint32_t dump_stride, const uint32_t *stride16, std::string &err) {
    vg_scan_params P; fill_params(L, S, stride16, P);
    P.dump = dump; P.dump_stride = dump_stride;
    CK(cudaMemsetAsync(S.d_counters, 0, kCounterWords * 4, L.stream));
    CK(launch_walk(fmt, mode, P, levels_[S.task.level].blocks, L.stream));
    launches++; h2d_bytes += sizeof(vg_scan_params);            // kernel parameters travel with the launch
    if (fmt == VG_FMT_P2TR) {
        // stage 1 wrote the internal keys; redo flagged batches on the device (count read there), then the tweak stage
        uint32_t k8[8]; S.task.base.to_limbs32(k8);
        CK(cudaMemcpyAsync(S.d_base8, k8, 32, cudaMemcpyHostToDevice, L.stream));
        CK(launch_slow(fmt, mode, P, S.d_base8, devices_[L.dev].d_pow2, 0xffffffffu, L.stream));
        if (p2tr_internal_) CK(launch_dump_points_x(P, L.stream));
        else CK(launch_taptweak(mode, P, L.stream));
        launches += 2; h2d_bytes += 32 + 2 * sizeof(vg_scan_params);
    }
    CK(cudaMemcpyAsync(S.h_counters, S.d_counters, kCounterWords * 4, cudaMemcpyDeviceToHost, L.stream));
    d2h_bytes += kCounterWords * 4;
    CK(cudaEventRecord(S.done, L.stream));
    S.busy = true;
    // after this launch the centres stand one stride further (meaningful for the final level only: plan.h)
    L.centers_base = S.task.base.add(U256(levels_[S.task.level].window).mul64(shard_count_ * (uint32_t)lanes_.size()));
    return true;
}

// ---- one scan ----
namespace {
struct WindowResult { uint64_t keys = 0; std::vector<std::future<HitBatch>> jobs; };
}
// hit records of one collected window (or of one piece of a window that overflowed its hit buffer), already split by key
// index into the buckets the confirm jobs take
struct Engine::Harvest {
    struct Piece { U256 base; std::vector<std::vector<vg_hit_rec>> buckets; };
    std::vector<Piece> pieces;
};
struct Engine::ScanRun {
    const WindowPlan *plan = nullptr;
    U256 start;
    int fmt = 0, first_level = 0, final_level = 0;
    uint32_t nl = 0, gl0 = 0;                 // lanes of this context, its first global lane
    std::mutex mu;
    std::condition_variable cv_main, cv_work;
    std::vector<LaneCursor> cur;              // per logical lane: next window of its sequence
    std::vector<char> done;                   // per logical lane: sequence exhausted
    std::vector<int> owner;                   // physical lane that runs logical lane i; -1: its device failed, anyone takes it
    std::vector<WindowTask> requeue;          // windows that were in flight on a device when it failed
    std::map<uint64_t, WindowResult> results;
    uint64_t merged = 0;                      // next window sequence number the calling thread merges
    uint64_t outstanding = 0;                 // windows issued and not yet delivered
    std::atomic<bool> stopping{false};
    int live_devices = 0, live_workers = 0;
    std::string fatal;
    int inject_dev = -1; long inject_at = -1;  // VGEN_FAULT_INJECT=<device index in the context>:<n>: its n-th window issue fails

    uint64_t seq_of(const LaneCursor &c, uint32_t li) const {
        const uint64_t ramp = (uint64_t)(final_level - first_level);
        return c.level < final_level ? (uint64_t)(c.level - first_level) * nl + li : (ramp + c.k) * nl + li;
    }
    // next window for physical lane p (R.mu held): 0 = task, 1 = nothing right now, 2 = nothing ever again
    int next_task(uint32_t p, WindowTask &t) {
        if (stopping.load()) return 2;
        int best_req = -1; int best_q = -1; uint64_t best = ~0ull;
        U256 best_off; uint64_t best_limit = 0;
        for (size_t i = 0; i < requeue.size(); i++) if (requeue[i].seq < best) { best = requeue[i].seq; best_req = (int)i; }
        bool all_done = true;
        for (uint32_t q = 0; q < nl; q++) {
            if (done[q]) continue;
            U256 off; uint64_t limit;
            if (!plan->window(gl0 + q, cur[q], off, limit)) { done[q] = 1; continue; }
            all_done = false;
            if (owner[q] != (int)p && owner[q] != -1) continue;
            uint64_t s = seq_of(cur[q], q);
            if (s < best) { best = s; best_req = -1; best_q = (int)q; best_off = off; best_limit = limit; }
        }
        if (best_req < 0 && best_q < 0) return (all_done && requeue.empty() && outstanding == 0) ? 2 : 1;
        if (best >= merged + kLookaheadPerLane * nl) return 1;          // far ahead of the merge: let it catch up first
        if (best_req >= 0) { t = requeue[best_req]; requeue.erase(requeue.begin() + best_req); }
        else {
            t.seq = best; t.logical = (uint32_t)best_q; t.base = start.add(best_off); t.limit = best_limit; t.level = cur[best_q].level;
            plan->advance(cur[best_q]);
        }
        outstanding++;
        return 0;
    }
};

// device records -> host buckets through the slot's pinned staging buffer
bool Engine::fetch_records(Lane &L, Slot &S, uint32_t n_hits, const U256 &base, uint64_t limit, Harvest &out, std::string &err) {
    Harvest::Piece pc; pc.base = base;
    // big windows are confirmed in several slices so that the last window's job is not a serial tail; records are scattered
    // by key index, so the slices are disjoint ascending key ranges and their sorted outputs concatenate in order
    const uint32_t n_slices = n_hits < 256 ? 1u : n_hits < (1u << 16) ? 4u : 16u;
    pc.buckets.resize(n_slices);
    for (auto &b : pc.buckets) b.reserve(n_hits / n_slices + n_hits / (4 * n_slices) + 16);
    const uint32_t stage = std::min(hit_cap_, kStageRecords);
    const uint64_t per = (limit + n_slices - 1) / n_slices;
    for (uint32_t done = 0; done < n_hits;) {
        uint32_t chunk = std::min(n_hits - done, stage);
        CK(cudaMemcpyAsync(S.h_hits, S.d_hits + done, (size_t)chunk * sizeof(vg_hit_rec), cudaMemcpyDeviceToHost, L.copy_stream));
        CK(cudaStreamSynchronize(L.copy_stream));
        d2h_bytes += (uint64_t)chunk * sizeof(vg_hit_rec);
        if (n_slices == 1) pc.buckets[0].insert(pc.buckets[0].end(), S.h_hits, S.h_hits + chunk);
        else for (uint32_t i = 0; i < chunk; i++) {
            const vg_hit_rec &r = S.h_hits[i];
            uint64_t idx = ((uint64_t)r.idx_hi << 32) | r.idx_lo;
            pc.buckets[std::min<uint64_t>(idx / per, n_slices - 1)].push_back(r);
        }
        done += chunk;
    }
    out.pieces.push_back(std::move(pc));
    return true;
}

// Collects the window in slot S (launched earlier): waits for it, redoes degenerate batches key by key, and moves the hit
// records to the host. A window with more hits than the hit buffer holds is re-run in pieces of the rescan level.
bool Engine::harvest(Lane &L, Slot &S, int fmt, ScanRun &R, Harvest &out, std::string &err) {
    (void)R;
    CK(cudaEventSynchronize(S.done));
    // degenerate batches of the window just finished + record fetch; used for the window itself and for rescan pieces
    auto finish = [&](uint32_t &n_hits) -> bool {
        if (S.h_counters[2]) {      // only a checked build (-DVG_CHECKED) ever writes here
            err = "device bounds check " + std::to_string(S.h_counters[2]) + " failed (" + std::to_string(S.h_counters[3]) + " times)";
            return false;
        }
        n_hits = S.h_counters[0];
        const uint32_t n_degen = S.h_counters[1];
        if (n_degen && fmt == VG_FMT_P2TR) {
            L.centers_level = -1;                       // already redone on the device between the two stages (launch_window)
        } else if (n_degen) {
            // batches that touched the point at infinity or collided with a table point: redo them key by key
            vg_scan_params P; fill_params(L, S, L.stride_pt[S.task.level].data(), P);
            uint32_t k8[8]; S.task.base.to_limbs32(k8);
            CK(cudaMemcpyAsync(S.d_base8, k8, 32, cudaMemcpyHostToDevice, L.stream));
            CK(launch_slow(fmt, VG_MODE_MATCH, P, S.d_base8, devices_[L.dev].d_pow2, std::min(n_degen, levels_[S.task.level].threads), L.stream));
            launches++;
            CK(cudaMemcpyAsync(S.h_counters, S.d_counters, kCounterWords * 4, cudaMemcpyDeviceToHost, L.stream));
            CK(cudaStreamSynchronize(L.stream));
            n_hits = S.h_counters[0];
            L.centers_level = -1;                       // centres of the flagged threads are void: set up again
        }
        return true;
    };
    uint32_t n_hits = 0;
    if (!finish(n_hits)) return false;
    if (n_hits <= hit_cap_) return fetch_records(L, S, n_hits, S.task.base, S.task.limit, out, err);
    // overflow: more hits than records. Re-run the window as consecutive windows of the rescan level, each of which has at
    // most hit_cap_ keys and therefore fits.
    overflow_rescans++;
    const WindowTask whole = S.task;
    const Level &RL = levels_[rescan_level_];
    bool ok = true;
    for (uint64_t off = 0; off < whole.limit && ok; off += RL.window) {
        S.task.level = rescan_level_; S.task.base = whole.base.add64(off); S.task.limit = std::min<uint64_t>(RL.window, whole.limit - off);
        ok = setup_centers(L, rescan_level_, S.task.base, err) &&
             launch_window(L, S, fmt, VG_MODE_MATCH, nullptr, 0, stride1_[rescan_level_].data(), err);
        if (ok && cudaEventSynchronize(S.done) != cudaSuccess) { err = "rescan window failed"; ok = false; }
        uint32_t n = 0;
        ok = ok && finish(n) && fetch_records(L, S, std::min(n, hit_cap_), S.task.base, S.task.limit, out, err);
    }
    S.task = whole;
    L.centers_level = -1;
    return ok;
}

void Engine::device_worker(ScanRun &R, int di) {
    Device &D = devices_[di];
    const int fmt = R.fmt;
    std::deque<std::pair<int, int>> inflight;          // (lane index, slot) in issue order
    long issued = 0;
    std::string err;
    // the device is lost: hand its unfinished windows and its lanes' future windows to the others
    auto fail = [&](const std::string &why, const WindowTask *extra) {
        std::lock_guard<std::mutex> g(R.mu);
        D.failed = true; D.failure = why;
        if (extra) { R.requeue.push_back(*extra); R.outstanding--; requeued_windows++; }
        for (auto &ps : inflight) {
            Slot &S = lanes_[ps.first].slot[ps.second];
            R.requeue.push_back(S.task); S.busy = false; R.outstanding--; requeued_windows++;
        }
        inflight.clear();
        for (int p : D.lanes) { if (R.owner[p] == p) R.owner[p] = -1; lanes_[p].centers_level = -1; }
        if (--R.live_devices == 0) { R.fatal = "device " + std::to_string(D.device) + " failed: " + why + " (no device left)"; R.stopping = true; }
        if (getenv("VGEN_TRACE")) fprintf(stderr, "[vgen trace] device %d dropped out: %s\n", D.device, why.c_str());
    };
    auto leave = [&]() {
        std::lock_guard<std::mutex> g(R.mu);
        R.live_workers--;
        R.cv_main.notify_all(); R.cv_work.notify_all();
    };
    if (cudaSetDevice(D.device) != cudaSuccess) { fail("cudaSetDevice failed", nullptr); leave(); return; }
    for (;;) {
        // issue: fill the free slots of this device's lanes, lowest window first
        bool finished = false;
        for (int round = 0; round < 2 && !finished; round++) for (int p : D.lanes) {
            Lane &L = lanes_[p];
            int s = !L.slot[0].busy ? 0 : (!L.slot[1].busy ? 1 : -1);
            if (s < 0) continue;
            WindowTask t; int rc;
            { std::lock_guard<std::mutex> g(R.mu); rc = R.next_task((uint32_t)p, t); }
            if (rc == 2) { finished = true; break; }
            if (rc == 1) continue;
            Slot &S = L.slot[s];
            S.task = t;
            bool ok = !(di == R.inject_dev && issued == R.inject_at);
            if (!ok) err = "injected fault (VGEN_FAULT_INJECT)";
            issued++;
            if (ok && (L.centers_level != t.level || !(L.centers_base == t.base))) ok = setup_centers(L, t.level, t.base, err);
            ok = ok && launch_window(L, S, fmt, VG_MODE_MATCH, nullptr, 0, L.stride_pt[t.level].data(), err);
            if (!ok) { S.busy = false; fail(err, &t); leave(); return; }
            inflight.push_back({p, s});
        }
        if (inflight.empty()) {
            if (finished) break;
            std::unique_lock<std::mutex> lk(R.mu);                       // nothing to do right now: the merge has to catch up, or
            R.cv_work.wait_for(lk, std::chrono::milliseconds(2));        // a failing device may still hand its windows over
            continue;
        }
        // collect the oldest window of this device
        auto [p, s] = inflight.front(); inflight.pop_front();
        Lane &L = lanes_[p]; Slot &S = L.slot[s];
        if (R.stopping.load()) {                                         // results no longer wanted: just wait for the launch
            cudaEventSynchronize(S.done);
            S.busy = false;
            std::lock_guard<std::mutex> g(R.mu);
            R.outstanding--;
            continue;
        }
        Harvest hv;
        if (!harvest(L, S, fmt, R, hv, err)) { S.busy = false; WindowTask t = S.task; fail(err, &t); leave(); return; }
        S.busy = false;
        WindowResult wr; wr.keys = S.task.limit;
        uint64_t n_rec = 0;
        for (auto &pc : hv.pieces) for (auto &b : pc.buckets) n_rec += b.size();
        const bool with_text = n_rec <= kEagerTextHits;
        const uint32_t dev_index = (uint32_t)L.device;
        for (auto &pc : hv.pieces) for (auto &b : pc.buckets) {
            if (b.empty()) continue;
            auto recs = std::make_shared<std::vector<vg_hit_rec>>(std::move(b));
            const U256 base = pc.base;
            wr.jobs.push_back(pool_->submit([this, recs, base, dev_index, fmt, with_text]() {
                // records in key order (the device appends them in atomic order)
                std::sort(recs->begin(), recs->end(), [](const vg_hit_rec &a, const vg_hit_rec &b) { return a.idx_hi != b.idx_hi ? a.idx_hi < b.idx_hi : a.idx_lo < b.idx_lo; });
                HitBatch out;
                out.hits.reserve(recs->size());
                if (with_text) out.text.reserve(recs->size() * 160);
                for (const vg_hit_rec &rec : *recs) {
                    Hit h; memset(&h, 0, sizeof h);
                    h.digest_len = (uint32_t)digest_len_of_format(fmt); h.lane = dev_index;
                    for (int w = 0; w < 8; w++) for (int b = 0; b < 4; b++) h.digest[4 * w + b] = (uint8_t)(rec.digest[w] >> (8 * b));
                    // exact confirmation on the host: re-encode the address from the digest and run the full DFA
                    std::string addr = format_address(fmt, h.digest);
                    if (!lp_.dfa.match(addr.data(), addr.size())) continue;
                    h.key = base.add64(((uint64_t)rec.idx_hi << 32) | rec.idx_lo);
                    if (with_text) {
                        uint8_t kb[32]; h.key.to_be(kb);
                        h.text_off = out.text.size();
                        out.text += h_hex(kb, 32); out.text += ' '; out.text += addr; out.text += ' '; out.text += format_wif(fmt, kb); out.text += '\n';
                        h.text_len = (uint32_t)(out.text.size() - h.text_off);
                    }
                    out.hits.push_back(h);
                }
                return out;
            }));
        }
        {
            std::lock_guard<std::mutex> g(R.mu);
            R.results.emplace(S.task.seq, std::move(wr));
            R.outstanding--;
        }
        R.cv_main.notify_all();
    }
    for (int p : D.lanes) cudaStreamSynchronize(lanes_[p].stream);
    leave();
}

bool Engine::scan(const U256 &start_in, const U256 *end_in, uint64_t max_hits, const volatile int32_t *stop, const Progress &progress,
                  std::vector<Hit> &hits, std::string &text, uint64_t &keys_scanned, std::string &err) {
    hits.clear(); text.clear(); keys_scanned = 0;
    if (!has_pattern_) { err = "no pattern set"; return false; }
    // valid scalars are 1..n-1 (src/address.rs:93); scalars outside are skipped, not counted (src/scanner.rs:294-295)
    U256 start = start_in, end = end_in ? *end_in : secp_order_minus1();
    if (end_in && end < start) { err = "range end is below range start"; return false; }
    if (start.is_zero()) start = U256(1);
    if (secp_order_minus1() < end) end = secp_order_minus1();
    if (end < start) return true;                      // nothing valid in range
    if (lp_.dev.kind == VG_MATCH_INTERVALS && lp_.dev.n_intervals == 0) {
        // the pattern can never match an address of this format: the reference would scan to the end of the range
        // and find nothing; report the range as scanned without launching
        U256 n = end.sub(start).add64(1);
        keys_scanned = n.fits64() ? n.w[0] : ~0ULL;
        return true;
    }
    if (n_failed_devices() == n_devices()) { err = "every device of this context has failed: " + devices_[0].failure; return false; }
    const U256 N = end.sub(start).add64(1);            // >= 1; cannot wrap because end <= n-1 < 2^256-1
    if (lp_.format == VG_FMT_P2TR && !ensure_taproot(N.fits64() ? N.w[0] : ~0ULL, err)) return false;
    // VGEN_TRACE=1: phase timestamps of this call on stderr (ms since entry)
    static const bool trace = getenv("VGEN_TRACE") != nullptr;
    const auto t_entry = std::chrono::steady_clock::now();
    auto mark = [&](const char *what) {
        if (trace) fprintf(stderr, "[vgen trace] %-22s %9.3f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_entry).count());
    };
    if (!compute_stride_points(err)) return false;
    mark("stride points");
    const uint32_t nl = (uint32_t)lanes_.size(), G = shard_count_ * nl;
    // ramp levels (small windows first) only when the scan is likely to end early: a hit limit, and a device matcher
    // dense enough that one full-size round would report a good part of that limit (e.g. generate -p ^1A -c 1)
    const double per_round = lp_.density * (double)levels_.back().window * (double)G;
    const bool ramp = max_hits != 0 && per_round > 0.25 * (double)max_hits;
    const WindowPlan plan(levels_, G, N, ramp);

    ScanRun R;
    R.plan = &plan; R.start = start; R.fmt = lp_.format; R.nl = nl; R.gl0 = shard_index_ * nl;
    R.first_level = plan.first_level(); R.final_level = plan.final_level();
    R.cur.assign(nl, plan.begin()); R.done.assign(nl, 0); R.owner.resize(nl);
    for (uint32_t q = 0; q < nl; q++) R.owner[q] = devices_[lanes_[q].dev].failed ? -1 : (int)q;
    if (const char *fi = getenv("VGEN_FAULT_INJECT")) {
        const char *colon = strchr(fi, ':');
        R.inject_dev = atoi(fi); R.inject_at = colon ? atol(colon + 1) : 0;
    }
    std::vector<std::thread> workers;
    for (int d = 0; d < (int)devices_.size(); d++) if (!devices_[d].failed) { R.live_devices++; R.live_workers++; }
    for (int d = 0; d < (int)devices_.size(); d++) if (!devices_[d].failed) workers.emplace_back([this, &R, d] { device_worker(R, d); });
    mark("workers started");

    // merge finished windows in sequence order on the calling thread (progress callback and stop flag live here)
    bool stopping = false;
    uint64_t n_merged = 0;
    {
        std::unique_lock<std::mutex> lk(R.mu);
        for (;;) {
            auto it = R.results.find(R.merged);
            if (it == R.results.end()) {
                if (R.live_workers == 0) break;
                R.cv_main.wait(lk);
                continue;
            }
            WindowResult wr = std::move(it->second);
            R.results.erase(it);
            R.merged++;
            lk.unlock();
            R.cv_work.notify_all();
            for (auto &f : wr.jobs) {
                HitBatch v = f.get();
                if (stopping || v.hits.empty()) continue;
                const uint64_t shift = text.size();
                const size_t at = hits.size();
                hits.insert(hits.end(), v.hits.begin(), v.hits.end());
                if (shift) for (size_t i = at; i < hits.size(); i++) hits[i].text_off += shift;
                text += v.text;
            }
            if (!stopping) {
                keys_scanned += wr.keys;
                if (trace && (n_merged++ < 3)) mark("window merged");
                if (progress) progress(keys_scanned);
                if (stop && *stop) stopping = true;
                if (max_hits && hits.size() >= max_hits) stopping = true;
                if (stopping) { R.stopping = true; R.cv_work.notify_all(); }
            }
            lk.lock();
        }
    }
    for (auto &t : workers) t.join();
    for (auto &kv : R.results) for (auto &f : kv.second.jobs) f.wait();     // windows finished after the stop: discarded
    mark("all windows merged");
    if (!R.fatal.empty()) { err = R.fatal; return false; }
    auto by_key = [](const Hit &a, const Hit &b) { return a.key < b.key; };
    if (!std::is_sorted(hits.begin(), hits.end(), by_key)) std::sort(hits.begin(), hits.end(), by_key);
    if (max_hits && hits.size() > max_hits) hits.resize(max_hits);
    mark("done");
    return true;
}

bool Engine::digest_batch(int format, const U256 &start, uint64_t n, uint8_t *out, uint32_t stride, std::string &err) {
    // format 6 (VGEN_P2TR_INTERNAL_KEY): the untweaked even-Y internal key, as the reference's await_result_p2tr returns it
    const bool internal = format == 6;
    if (internal) format = VG_FMT_P2TR;
    if (format < 0 || format > 5 || stride == 0 || stride > 32) { err = "invalid format or stride"; return false; }
    if (n == 0) return true;
    struct Flag { bool &f; ~Flag() { f = false; } } flag_guard{p2tr_internal_};
    p2tr_internal_ = internal;
    if (n > (1ull << 28)) { err = "digest_batch is limited to 2^28 keys per call"; return false; }
    memset(out, 0, (size_t)n * stride);
    if (!(start <= secp_order_minus1())) return true;           // every scalar invalid: all zeros
    {   // scalars >= n are invalid (zeros): only walk up to n-1
        U256 room = secp_order_minus1().sub(start).add64(1);
        if (room.fits64() && room.w[0] < n) n = room.w[0];
    }
    if (format == VG_FMT_P2TR && !ensure_taproot(n, err)) return false;
    Lane *Lp = nullptr;
    for (auto &L : lanes_) if (!devices_[L.dev].failed) { Lp = &L; break; }
    if (!Lp) { err = "every device of this context has failed"; return false; }
    Lane &L0 = *Lp;
    CK(cudaSetDevice(L0.device));
    uint8_t *d_out = nullptr;
    if (cudaMalloc(&d_out, (size_t)n * stride) != cudaSuccess) { err = "cudaMalloc failed for the digest buffer"; return false; }
    cudaMemsetAsync(d_out, 0, (size_t)n * stride, L0.stream);
    // pick the smallest level whose window covers n, else the full level; the lane walks alone with stride W
    int level = (int)levels_.size() - 1;
    for (int lv = 0; lv < (int)levels_.size(); lv++) if (levels_[lv].window >= n) { level = lv; break; }
    const uint64_t W = levels_[level].window;
    LoweredPattern saved = lp_;
    memset(&lp_.dev, 0, sizeof lp_.dev);
    bool good = true;
    U256 base = start;
    Slot &S = L0.slot[0];
    for (uint64_t off = 0; off < n && good; off += W) {
        S.task.base = base; S.task.level = level; S.task.limit = std::min<uint64_t>(W, n - off);
        if (off == 0) good = setup_centers(L0, level, base, err);
        if (good) good = launch_window(L0, S, format, VG_MODE_DUMP, d_out + off * stride, stride, stride1_[level].data(), err);
        if (good && cudaStreamSynchronize(L0.stream) != cudaSuccess) { err = "walk kernel failed"; good = false; }
        S.busy = false;
        if (good && S.h_counters[1] && format == VG_FMT_P2TR) {
            if (off + W < n) good = setup_centers(L0, level, base.add64(W), err);      // flagged batches were redone on the device
        } else if (good && S.h_counters[1]) {
            vg_scan_params P; fill_params(L0, S, stride1_[level].data(), P);
            P.dump = d_out + off * stride; P.dump_stride = stride;
            uint32_t k8[8]; base.to_limbs32(k8);
            cudaMemcpyAsync(S.d_base8, k8, 32, cudaMemcpyHostToDevice, L0.stream);
            cudaError_t e = launch_slow(format, VG_MODE_DUMP, P, S.d_base8, devices_[L0.dev].d_pow2, std::min(S.h_counters[1], levels_[level].threads), L0.stream);
            launches++;
            if (e != cudaSuccess || cudaStreamSynchronize(L0.stream) != cudaSuccess) { err = "slow kernel failed"; good = false; }
            if (good && off + W < n) good = setup_centers(L0, level, base.add64(W), err);
        }
        base = base.add64(W);
    }
    if (good && cudaMemcpy(out, d_out, (size_t)n * stride, cudaMemcpyDeviceToHost) != cudaSuccess) { err = "copy of digests failed"; good = false; }
    cudaFree(d_out);
    L0.centers_level = -1;
    lp_ = saved;
    return good;
}

bool Engine::bench_resident(const U256 &start, uint32_t n_windows, float &ms, uint64_t &keys, uint64_t &nlaunch, std::string &err) {
    if (!has_pattern_) { err = "no pattern set"; return false; }
    if (lp_.format == VG_FMT_P2TR && !ensure_taproot(~0ULL, err)) return false;      // sustained rate: the long-job table
    if (!compute_stride_points(err)) return false;
    // every device of the context at once, its lanes exactly as a scan drives them: window k of the sequence goes to global
    // lane k mod G; n_windows windows per device
    const int fl = (int)levels_.size() - 1;
    const uint64_t W = levels_[fl].window;
    const uint32_t nl = (uint32_t)lanes_.size();
    std::vector<vg_scan_params> Ps(nl);
    for (uint32_t j = 0; j < nl; j++) {
        Lane &L = lanes_[j];
        if (devices_[L.dev].failed) { err = "a device of this context has failed"; return false; }
        CK(cudaSetDevice(L.device));
        uint64_t gl = (uint64_t)shard_index_ * nl + j;
        U256 base = start.add(U256(gl).mul64(W));
        if (!setup_centers(L, fl, base, err)) return false;
        Slot &S = L.slot[0];
        S.task.base = base; S.task.level = fl; S.task.limit = W;
        fill_params(L, S, L.stride_pt[fl].data(), Ps[j]);
        CK(cudaStreamSynchronize(L.stream));
    }
    const size_t nd = devices_.size();
    std::vector<cudaEvent_t> e0(nd), e1(nd);
    std::vector<std::vector<cudaEvent_t>> done(nd);
    for (size_t d = 0; d < nd; d++) {
        Device &D = devices_[d];
        CK(cudaSetDevice(D.device));
        CK(cudaEventCreate(&e0[d])); CK(cudaEventCreate(&e1[d]));
        done[d].resize(D.lanes.size());
        for (auto &ev : done[d]) CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        cudaStream_t s0 = lanes_[D.lanes[0]].stream;
        CK(cudaEventRecord(e0[d], s0));
        for (size_t j = 1; j < D.lanes.size(); j++) CK(cudaStreamWaitEvent(lanes_[D.lanes[j]].stream, e0[d], 0));      // no lane starts before e0
    }
    for (uint32_t w = 0; w < n_windows; w++) for (size_t d = 0; d < nd; d++) {
        Device &D = devices_[d];
        const int li = D.lanes[w % D.lanes.size()];
        Lane &L = lanes_[li]; Slot &S = L.slot[0];
        CK(cudaSetDevice(D.device));
        CK(cudaMemsetAsync(S.d_counters, 0, kCounterWords * 4, L.stream));
        CK(launch_walk(lp_.format, VG_MODE_MATCH, Ps[li], levels_[fl].blocks, L.stream));
        launches++;
        if (lp_.format == VG_FMT_P2TR) {
            // the lane's (w / lanes)-th window: its centres have moved on by that many strides of G lanes
            const uint64_t G = (uint64_t)shard_count_ * nl;
            S.task.base.add(U256(W).mul64(G).mul64(w / D.lanes.size())).to_limbs32(Ps[li].tap_base);
            CK(launch_taptweak(VG_MODE_MATCH, Ps[li], L.stream)); launches++;
        }
    }
    ms = 0;
    for (size_t d = 0; d < nd; d++) {
        Device &D = devices_[d];
        CK(cudaSetDevice(D.device));
        cudaStream_t s0 = lanes_[D.lanes[0]].stream;
        for (size_t j = 1; j < D.lanes.size(); j++) { CK(cudaEventRecord(done[d][j], lanes_[D.lanes[j]].stream)); CK(cudaStreamWaitEvent(s0, done[d][j], 0)); }
        CK(cudaEventRecord(e1[d], s0));                                                       // after every lane's last window
    }
    for (size_t d = 0; d < nd; d++) {
        CK(cudaSetDevice(devices_[d].device));
        CK(cudaEventSynchronize(e1[d]));
        float t = 0;
        CK(cudaEventElapsedTime(&t, e0[d], e1[d]));
        ms = std::max(ms, t);
        cudaEventDestroy(e0[d]); cudaEventDestroy(e1[d]);
        for (auto &ev : done[d]) cudaEventDestroy(ev);
    }
    nlaunch = (uint64_t)n_windows * nd;
    keys = (uint64_t)n_windows * nd * W;
    for (auto &L : lanes_) L.centers_level = -1;
    return true;
}

}  // namespace vg
