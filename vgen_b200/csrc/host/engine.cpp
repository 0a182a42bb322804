// Host driver of the B200 scan path (see engine.h).
//
// Window schedule. With G global lanes (shards x GPUs) the keyspace [start, start+N) is cut into a fixed
// sequence of windows: one small window per lane at each ramp level (so short ranges and dense patterns
// finish without computing a full-size window), then full-size windows handed out round-robin: lane g
// takes windows g, g+G, g+2G, ... Each lane therefore walks with a disjoint start scalar and a common
// stride of G*W keys (the walk kernel adds the stride point (G*W)*G to every centre), there is no
// device-to-device traffic, and hits come back in ascending key order.
#include "engine.h"
#include <algorithm>
#include <chrono>
#include <deque>
#include <future>
#include <string.h>
#include <stdio.h>
#include <stdlib.h>

namespace vg {

#define CK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { err = std::string(#call) + ": " + cudaGetErrorString(e_); return false; } } while (0)

static const int kCounterWords = 4;
static const uint32_t kEagerTextHits = 1u << 16;   // windows with at most this many device hits get their text records built by the confirm job

Engine::~Engine() { destroy(); }

int Engine::lanes_per_device() {
    const char *e = getenv("VGEN_LANES_PER_GPU");       // tuning override
    int n = e ? atoi(e) : 2;
    return n < 1 ? 1 : (n > 4 ? 4 : n);
}

void Engine::destroy() {
    for (auto &L : lanes_) {
        cudaSetDevice(L.device);
        if (L.stream) cudaStreamSynchronize(L.stream);
        for (auto &S : L.slot) {
            if (S.d_hits) cudaFree(S.d_hits);
            if (S.d_counters) cudaFree(S.d_counters);
            if (S.d_degen) cudaFree(S.d_degen);
            if (S.d_base8) cudaFree(S.d_base8);
            if (S.h_counters) cudaFreeHost(S.h_counters);
            if (S.done) cudaEventDestroy(S.done);
        }
        for (void *p : {(void *)L.d_pow2, (void *)L.d_gtab, (void *)L.d_centers, (void *)L.d_scratch, (void *)L.d_base8,
                        (void *)L.d_stride_pts, (void *)L.d_points, (void *)(L.owns_tap_table ? L.d_tap_table : nullptr), (void *)L.d_dfa_table, (void *)L.d_dfa_accept})
            if (p) cudaFree(p);
        if (L.h_hits) cudaFreeHost(L.h_hits);
        if (L.own_stream && L.stream) cudaStreamDestroy(L.stream);
        if (L.copy_stream) cudaStreamDestroy(L.copy_stream);
    }
    lanes_.clear();
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
    int min_sm = 1 << 30;
    for (int d : devs) {
        cudaDeviceProp prop;
        CK(cudaGetDeviceProperties(&prop, d));
        if (prop.major < 10) { err = std::string("device ") + prop.name + " is not sm_100 class; this library is built for B200 (sm_100a) only"; return false; }
        Lane L; L.device = d; L.sm_count = prop.multiProcessorCount; L.name = prop.name;
        min_sm = std::min(min_sm, L.sm_count);
        // Several lanes (stream + centres + scratch + hit buffers) per GPU: a window's warps do not finish together
        // (ncu r01: 12.9 of 16 resident warps active on average), and a second stream's next window fills the CTA slots the
        // early finishers free. Measured on B200: 1 lane 5,819, 2 lanes 6,094, 3 lanes 6,085 Mkeys/s (profiles/overlap_r01.log).
        for (int k = 0; k < lanes_per_device(); k++) lanes_.push_back(L);
    }
    levels_ = compute_levels(min_sm, hint, half_, VG_BLOCK, VG_CTAS_PER_SM);
    const Level &F = levels_.back();
    for (auto &L : lanes_) {
        CK(cudaSetDevice(L.device));
        CK(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
        CK(cudaStreamCreateWithFlags(&L.copy_stream, cudaStreamNonBlocking));
        CK(cudaMalloc(&L.d_pow2, (size_t)VG_GTABLE_WORDS * 4));       // pow2[256] + win8[32][256] (ec.cuh)
        CK(cudaMalloc(&L.d_gtab, (size_t)half_ * 16 * 4));
        CK(cudaMalloc(&L.d_centers, (size_t)16 * F.threads * 4));
        CK(cudaMalloc(&L.d_scratch, (size_t)(half_ + 1) * 8 * F.threads * 4));
        CK(cudaMalloc(&L.d_base8, 64));
        CK(cudaMalloc(&L.d_stride_pts, 16 * 4 * 8));
        for (auto &S : L.slot) {
            CK(cudaMalloc(&S.d_hits, (size_t)F.window * sizeof(vg_hit_rec)));
            CK(cudaMalloc(&S.d_counters, kCounterWords * 4));
            CK(cudaMalloc(&S.d_degen, (size_t)F.threads * 4));
            CK(cudaMalloc(&S.d_base8, 32));
            CK(cudaHostAlloc(&S.h_counters, kCounterWords * 4, cudaHostAllocDefault));
            CK(cudaEventCreateWithFlags(&S.done, cudaEventDisableTiming));
        }
        L.h_hits_cap = 1u << 20;
        CK(cudaHostAlloc(&L.h_hits, L.h_hits_cap * sizeof(vg_hit_rec), cudaHostAllocDefault));
        CK(launch_build_tables(L.d_pow2, L.d_gtab, half_, L.stream));
        launches += 3;
        CK(cudaStreamSynchronize(L.stream));
    }
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
    for (auto &L : lanes_) {
        CK(cudaSetDevice(L.device));
        if (L.d_dfa_table) { cudaFree(L.d_dfa_table); L.d_dfa_table = nullptr; }
        if (L.d_dfa_accept) { cudaFree(L.d_dfa_accept); L.d_dfa_accept = nullptr; }
        CK(cudaMalloc(&L.d_dfa_table, d.table.size() * 2 + 4));     // the kernel stages these word-wise
        CK(cudaMalloc(&L.d_dfa_accept, d.accept_end.size() + 4));
        CK(cudaMemcpy(L.d_dfa_table, d.table.data(), d.table.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(L.d_dfa_accept, d.accept_end.data(), d.accept_end.size(), cudaMemcpyHostToDevice));
    }
    return true;
}

// Window width of the P2TR fixed-base table (taproot.cuh), chosen per job (measured on B200, profiles/tap_bits_r01.log):
//   16 bits: 16 windows x 65,536 x 64 B = 64 MiB, built in 2 ms; 16 additions per key                      476 Mkeys/s
//   24 bits: 11 windows x 2^24 x 64 B = 11 GiB of the 180 GB HBM, built in 0.15 s; 11 additions per key     638 Mkeys/s
//   26 bits: 10 windows x 2^26 x 64 B = 40 GiB, built in 0.43 s; 10 additions per key                       683 Mkeys/s
// The entries are gathered at random from DRAM either way (ncu: the 64 MiB table is flushed out of L2 by the window's point
// stream), so a wider table only saves additions. Jobs of >= 2^30 keys take 24 bits, >= 2^36 keys 26 bits (when the
// device has the room). VGEN_TAP_BITS (4..26) forces a width.
static uint32_t tap_bits_for(uint64_t keys_hint) {
    static const char *e = getenv("VGEN_TAP_BITS");
    if (e) { long v = strtol(e, nullptr, 10); return (uint32_t)(v < 4 ? 4 : v > 26 ? 26 : v); }
    return keys_hint >= (1ull << 36) ? 26u : keys_hint >= (1ull << 30) ? 24u : 16u;
}
static size_t tap_table_bytes(uint32_t bits) { return ((size_t)((256 + bits - 1) / bits) << bits) * 64; }

// P2TR needs two more buffers per lane: the stage-1 -> stage-2 point buffer (64 B per window key) and the
// fixed-base table of the tweak addition (one per GPU, shared by its lanes). Allocated and built on first use; a
// wider table replaces a narrower one when a big job comes, never the other way round.
bool Engine::ensure_taproot(uint64_t keys_hint, std::string &err) {
    const Level &F = levels_.back();
    uint32_t want = tap_bits_for(keys_hint);
    for (auto &L : lanes_) {
        if (L.d_points) continue;
        CK(cudaSetDevice(L.device));
        CK(cudaMalloc(&L.d_points, (size_t)F.window * 64));
    }
    if (tap_bits_ != 0 && (tap_bits_ >= want || getenv("VGEN_TAP_BITS"))) return true;
    for (auto &L : lanes_) {            // a big table needs room: fall back to the small one on a crowded device
        size_t free_b = 0, total_b = 0;
        CK(cudaSetDevice(L.device));
        CK(cudaMemGetInfo(&free_b, &total_b));
        const size_t held = tap_bits_ ? tap_table_bytes(tap_bits_) : 0;      // freed before the new one is allocated
        if (want > 24 && free_b + held < tap_table_bytes(want) + (8ull << 30)) want = 24;
        if (want > 16 && free_b + held < tap_table_bytes(want) + (4ull << 30)) want = 16;
    }
    if (tap_bits_ >= want && tap_bits_ != 0) return true;
    const auto t0 = std::chrono::steady_clock::now();
    for (auto &L : lanes_) { CK(cudaSetDevice(L.device)); CK(cudaStreamSynchronize(L.stream)); }
    for (auto &L : lanes_) {
        if (L.owns_tap_table && L.d_tap_table) { CK(cudaSetDevice(L.device)); CK(cudaFree(L.d_tap_table)); }
        L.d_tap_table = nullptr; L.owns_tap_table = false;
    }
    for (auto &L : lanes_) {
        for (auto &O : lanes_) if (&O != &L && O.device == L.device && O.d_tap_table) L.d_tap_table = O.d_tap_table;
        if (L.d_tap_table) continue;
        CK(cudaSetDevice(L.device));
        CK(cudaMalloc(&L.d_tap_table, tap_table_bytes(want)));
        L.owns_tap_table = true;
        CK(launch_build_tap_table(L.d_tap_table, want, L.d_pow2, L.stream));
        launches++;
    }
    for (auto &L : lanes_) { CK(cudaSetDevice(L.device)); CK(cudaStreamSynchronize(L.stream)); }
    tap_bits_ = want;
    if (getenv("VGEN_TRACE"))
        fprintf(stderr, "[vgen trace] tap table %u bits     %9.3f ms\n", want, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
    return true;
}

bool Engine::set_pattern(int format, const std::string &pattern, bool ci, std::string &err) {
    LoweredPattern lp;
    if (!lower_pattern(format, pattern, ci, lp, err)) return false;
    lp_ = lp;
    has_pattern_ = true;
    return upload_dfa(err);
}

// stride point of the final level for the current global lane count: (G * W_final) * G
bool Engine::compute_stride_points(std::string &err) {
    uint32_t G = shard_count_ * (uint32_t)lanes_.size();
    if (stride_lanes_ == G) return true;
    for (auto &L : lanes_) {
        CK(cudaSetDevice(L.device));
        L.stride_pt.assign(levels_.size(), {});
        for (size_t lv = 0; lv < levels_.size(); lv++) {
            // every level uses its own G*W stride; only the final level's next centre is ever consumed
            U256 k = U256(levels_[lv].window).mul64(G);
            uint32_t k8[8]; k.to_limbs32(k8);
            CK(cudaMemcpyAsync(L.d_base8, k8, 32, cudaMemcpyHostToDevice, L.stream));
            CK(launch_points_of_scalars(L.d_stride_pts, 1, L.d_base8, 0, L.d_pow2, L.stream));
            launches++;
            CK(cudaMemcpyAsync(L.stride_pt[lv].data(), L.d_stride_pts, 64, cudaMemcpyDeviceToHost, L.stream));
            CK(cudaStreamSynchronize(L.stream));
        }
        L.centers_level = -1;
    }
    stride_lanes_ = G;
    return true;
}

bool Engine::setup_centers(Lane &L, int level, const U256 &window_base, std::string &err) {
    const Level &V = levels_[level];
    U256 first = window_base.add64(half_);
    uint32_t k8[8]; first.to_limbs32(k8);
    CK(cudaMemcpyAsync(L.d_base8, k8, 32, cudaMemcpyHostToDevice, L.stream));
    CK(launch_points_of_scalars(L.d_centers, V.threads, L.d_base8, 2ull * half_, L.d_pow2, L.stream));
    launches++; h2d_bytes += 32;
    L.centers_level = level;
    return true;
}

void Engine::fill_params(Lane &L, Slot &S, vg_scan_params &P) {
    const Level &V = levels_[S.level];
    memset(&P, 0, sizeof P);
    P.n_threads = V.threads; P.half = half_;
    P.centers = L.d_centers; P.scratch = L.d_scratch; P.gtab = L.d_gtab;
    memcpy(P.stride_pt, L.stride_pt[S.level].data(), 64);
    P.limit = S.limit;
    P.hits = S.d_hits; P.hit_count = S.d_counters; P.hit_cap = (uint32_t)std::min<uint64_t>(V.window, 0xffffffffu);
    P.degen_list = S.d_degen; P.degen_count = S.d_counters + 1; P.degen_cap = V.threads;
    P.points = L.d_points; P.tap_table = L.d_tap_table; P.tap_bits = tap_bits_;
    P.m = lp_.dev;
    P.m.dfa.table = L.d_dfa_table; P.m.dfa.accept = L.d_dfa_accept;
    P.dfa_smem_bytes = (P.m.kind == VG_MATCH_DFA || P.m.confirm_dfa) ? dfa_smem_bytes(P.m.dfa) : 0;
}

bool Engine::launch_window(Lane &L, int s, int fmt, int mode, uint8_t *dump, uint32_t dump_stride, std::string &err) {
    Slot &S = L.slot[s];
    vg_scan_params P; fill_params(L, S, P);
    P.dump = dump; P.dump_stride = dump_stride;
    CK(cudaMemsetAsync(S.d_counters, 0, kCounterWords * 4, L.stream));
    CK(launch_walk(fmt, mode, P, levels_[S.level].blocks, L.stream));
    launches++; h2d_bytes += sizeof(vg_scan_params);            // kernel parameters travel with the launch
    if (fmt == VG_FMT_P2TR) {
        // stage 1 wrote the internal keys; redo flagged batches on the device (count read there), then the tweak stage
        uint32_t k8[8]; S.base.to_limbs32(k8);
        CK(cudaMemcpyAsync(S.d_base8, k8, 32, cudaMemcpyHostToDevice, L.stream));
        CK(launch_slow(fmt, mode, P, S.d_base8, L.d_pow2, 0xffffffffu, L.stream));
        if (p2tr_internal_) CK(launch_dump_points_x(P, L.stream));
        else CK(launch_taptweak(mode, P, L.stream));
        launches += 2; h2d_bytes += 32 + 2 * sizeof(vg_scan_params);
    }
    CK(cudaMemcpyAsync(S.h_counters, S.d_counters, kCounterWords * 4, cudaMemcpyDeviceToHost, L.stream));
    d2h_bytes += kCounterWords * 4;
    CK(cudaEventRecord(S.done, L.stream));
    S.busy = true;
    return true;
}

namespace {
struct LaneRun { LaneCursor cur; bool done = false; U256 expect_base; bool have_expect = false; };
}

bool Engine::scan(const U256 &start_in, const U256 *end_in, uint64_t max_hits, const volatile int32_t *stop, const Progress &progress,
                  std::vector<Hit> &hits, uint64_t &keys_scanned, std::string &err) {
    hits.clear(); keys_scanned = 0;
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
    const int fl = (int)levels_.size() - 1;
    // ramp levels (small windows first) only when the scan is likely to end early: a hit limit, and a device matcher
    // dense enough that one full-size round would report a good part of that limit (e.g. generate -p ^1A -c 1)
    const double per_round = lp_.density * (double)levels_.back().window * (double)G;
    const bool ramp = max_hits != 0 && per_round > 0.25 * (double)max_hits;
    const WindowPlan plan(levels_, G, N, ramp);
    std::vector<LaneRun> runs(nl);
    for (auto &r : runs) r.cur = plan.begin();
    uint64_t seq = 0;
    const int fmt = lp_.format;

    std::deque<std::pair<uint32_t, int>> inflight;
    auto issue = [&](uint32_t li) -> bool {
        Lane &L = lanes_[li]; LaneRun &r = runs[li];
        if (r.done) return true;
        int s = !L.slot[0].busy ? 0 : (!L.slot[1].busy ? 1 : -1);
        if (s < 0) return true;
        U256 off; uint64_t limit;
        if (!plan.window(shard_index_ * nl + li, r.cur, off, limit)) { r.done = true; return true; }
        if (cudaSetDevice(L.device) != cudaSuccess) { err = "cudaSetDevice failed"; return false; }
        Slot &S = L.slot[s];
        const int level = r.cur.level;
        S.base = start.add(off); S.level = level; S.seq = seq++; S.limit = limit;
        if (L.centers_level != level || !r.have_expect || !(r.expect_base == S.base)) {
            if (!setup_centers(L, level, S.base, err)) return false;
        }
        if (!launch_window(L, s, fmt, VG_MODE_MATCH, nullptr, 0, err)) return false;
        // after this launch the centres stand at base + G*W of the same level
        r.expect_base = S.base.add(plan.stride(level)); r.have_expect = level == fl;
        plan.advance(r.cur);
        inflight.push_back({li, s});
        return true;
    };
    std::deque<std::future<std::vector<Hit>>> pending;   // per-window confirm jobs, in window order
    auto drain = [&](bool only_ready) {
        while (!pending.empty()) {
            if (only_ready && pending.front().wait_for(std::chrono::seconds(0)) == std::future_status::timeout) break;
            std::vector<Hit> v = pending.front().get();
            pending.pop_front();
            for (auto &h : v) hits.push_back(std::move(h));
        }
    };
    auto collect = [&](uint32_t li, int s, bool keep) -> bool {
        Lane &L = lanes_[li]; Slot &S = L.slot[s];
        CK(cudaSetDevice(L.device));
        CK(cudaEventSynchronize(S.done));
        S.busy = false;
        if (!keep) return true;
        uint32_t n_hits = S.h_counters[0], n_degen = S.h_counters[1];
        if (n_degen && fmt == VG_FMT_P2TR) {
            // already redone on the device between the two stages (launch_window); only the centres need a new set-up
            L.centers_level = -1;
            runs[li].have_expect = false;
        } else if (n_degen) {
            // batches that touched the point at infinity or collided with a table point: redo them key by key
            vg_scan_params P; fill_params(L, S, P);
            uint32_t k8[8]; S.base.to_limbs32(k8);
            CK(cudaStreamSynchronize(L.stream));          // the other slot's window may still be running; d_base8 is shared
            bool other_busy = L.slot[1 - s].busy;
            CK(cudaMemcpyAsync(L.d_base8, k8, 32, cudaMemcpyHostToDevice, L.stream));
            CK(launch_slow(fmt, VG_MODE_MATCH, P, L.d_base8, L.d_pow2, std::min(n_degen, levels_[S.level].threads), L.stream));
            launches++;
            CK(cudaMemcpyAsync(S.h_counters, S.d_counters, kCounterWords * 4, cudaMemcpyDeviceToHost, L.stream));
            CK(cudaStreamSynchronize(L.stream));
            n_hits = S.h_counters[0];
            L.centers_level = -1;                         // centres of the flagged threads are void: set up again
            runs[li].have_expect = false;
            (void)other_busy;
        }
        uint64_t cap = levels_[S.level].window;
        if (n_hits > cap) n_hits = (uint32_t)cap;         // cannot happen: at most one hit per key
        // device records -> host vector (pinned staging is reused by the next window), then confirm (+ format when the
        // window is sparse) on a worker thread so that it overlaps the windows still running on the GPU
        auto recs = std::make_shared<std::vector<vg_hit_rec>>();
        recs->reserve(n_hits);
        for (uint32_t done = 0; done < n_hits;) {
            uint32_t chunk = (uint32_t)std::min<size_t>(n_hits - done, L.h_hits_cap);
            CK(cudaMemcpyAsync(L.h_hits, S.d_hits + done, (size_t)chunk * sizeof(vg_hit_rec), cudaMemcpyDeviceToHost, L.copy_stream));
            CK(cudaStreamSynchronize(L.copy_stream));
            d2h_bytes += (uint64_t)chunk * sizeof(vg_hit_rec);
            recs->insert(recs->end(), L.h_hits, L.h_hits + chunk);
            done += chunk;
        }
        const U256 base = S.base;
        const bool with_text = n_hits <= kEagerTextHits;
        // records in key order (the device appends them in atomic order): slices then come out ascending and consecutive, so the
        // hit list stays sorted as windows are collected and the final sort is a no-op check
        std::sort(recs->begin(), recs->end(), [](const vg_hit_rec &a, const vg_hit_rec &b) { return a.idx_hi != b.idx_hi ? a.idx_hi < b.idx_hi : a.idx_lo < b.idx_lo; });
        // big windows are confirmed in up to 4 slices so that the last window's job is not a serial tail
        const size_t n_rec = recs->size();
        const size_t n_slices = n_rec >= 256 ? 4 : 1;
        for (size_t sl = 0; sl < n_slices; sl++) {
            const size_t lo = n_rec * sl / n_slices, hi = n_rec * (sl + 1) / n_slices;
            const uint32_t dev_index = (uint32_t)L.device;
            auto job = [this, recs, base, dev_index, fmt, with_text, lo, hi]() {
                std::vector<Hit> out;
                out.reserve(hi - lo);
                for (size_t r = lo; r < hi; r++) {
                    const vg_hit_rec &R = (*recs)[r];
                    Hit h; memset(h.digest, 0, 32);
                    h.digest_len = (uint32_t)digest_len_of_format(fmt); h.lane = dev_index;
                    for (int w = 0; w < 8; w++) for (int b = 0; b < 4; b++) h.digest[4 * w + b] = (uint8_t)(R.digest[w] >> (8 * b));
                    // exact confirmation on the host: re-encode the address from the digest and run the full DFA
                    std::string addr = format_address(fmt, h.digest);
                    if (!lp_.dfa.match(addr.data(), addr.size())) continue;
                    h.key = base.add64(((uint64_t)R.idx_hi << 32) | R.idx_lo);
                    if (with_text) {
                        uint8_t kb[32]; h.key.to_be(kb);
                        h.text = h_hex(kb, 32) + " " + addr + " " + format_wif(fmt, kb) + "\n";
                    }
                    out.push_back(std::move(h));
                }
                return out;
            };
            pending.push_back(std::async(n_rec >= 64 ? std::launch::async : std::launch::deferred, job));
        }
        if (max_hits) drain(false);                       // a hit limit needs the confirmed count now
        else drain(true);                                  // otherwise only reap the jobs that are already finished
        keys_scanned += S.limit;
        return true;
    };

    bool stopping = false;
    for (int round = 0; round < 2; round++) for (uint32_t li = 0; li < nl; li++) if (!issue(li)) return false;
    mark("first windows issued");
    uint64_t n_collected = 0;
    while (!inflight.empty()) {
        auto [li, s] = inflight.front(); inflight.pop_front();
        if (!collect(li, s, !stopping)) return false;
        if (trace && (n_collected++ < 3 || inflight.size() < 3)) mark("window collected");
        if (!stopping) {
            if (progress) progress(keys_scanned);
            if (stop && *stop) stopping = true;
            if (max_hits && hits.size() >= max_hits) stopping = true;
        }
        if (!stopping && !issue(li)) return false;
    }
    mark("all windows collected");
    drain(false);
    mark("confirm jobs drained");
    for (auto &L : lanes_) { CK(cudaSetDevice(L.device)); CK(cudaStreamSynchronize(L.stream)); L.centers_level = -1; }
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
    Lane &L = lanes_[0];
    CK(cudaSetDevice(L.device));
    uint32_t saved_idx = shard_index_, saved_cnt = shard_count_;
    shard_index_ = 0; shard_count_ = 1; stride_lanes_ = 0;
    // single lane, stride = one window: emulate G = 1 by computing stride points for one lane only
    std::vector<Lane> others;
    if (lanes_.size() > 1) { others.assign(lanes_.begin() + 1, lanes_.end()); lanes_.resize(1); }
    bool ok = compute_stride_points(err);
    uint8_t *d_out = nullptr;
    auto restore = [&]() {
        if (d_out) cudaFree(d_out);
        for (auto &o : others) lanes_.push_back(o);
        shard_index_ = saved_idx; shard_count_ = saved_cnt; stride_lanes_ = 0;
    };
    if (!ok) { restore(); return false; }
    Lane &L0 = lanes_[0];
    if (cudaMalloc(&d_out, (size_t)n * stride) != cudaSuccess) { err = "cudaMalloc failed for the digest buffer"; restore(); return false; }
    cudaMemsetAsync(d_out, 0, (size_t)n * stride, L0.stream);
    // pick the smallest level whose window covers n, else the full level
    int level = (int)levels_.size() - 1;
    for (int lv = 0; lv < (int)levels_.size(); lv++) if (levels_[lv].window >= n) { level = lv; break; }
    const uint64_t W = levels_[level].window;
    LoweredPattern saved = lp_;
    memset(&lp_.dev, 0, sizeof lp_.dev);
    bool good = true;
    U256 base = start;
    for (uint64_t off = 0; off < n && good; off += W) {
        Slot &S = L0.slot[0];
        S.base = base; S.level = level; S.limit = std::min<uint64_t>(W, n - off);
        if (off == 0) good = setup_centers(L0, level, base, err);
        if (good) good = launch_window(L0, 0, format, VG_MODE_DUMP, d_out + off * stride, stride, err);
        if (good && cudaStreamSynchronize(L0.stream) != cudaSuccess) { err = "walk kernel failed"; good = false; }
        S.busy = false;
        if (good && S.h_counters[1] && format == VG_FMT_P2TR) {
            if (off + W < n) good = setup_centers(L0, level, base.add64(W), err);      // flagged batches were redone on the device
        } else if (good && S.h_counters[1]) {
            vg_scan_params P; fill_params(L0, S, P);
            P.dump = d_out + off * stride; P.dump_stride = stride;
            uint32_t k8[8]; base.to_limbs32(k8);
            cudaMemcpyAsync(L0.d_base8, k8, 32, cudaMemcpyHostToDevice, L0.stream);
            cudaError_t e = launch_slow(format, VG_MODE_DUMP, P, L0.d_base8, L0.d_pow2, std::min(S.h_counters[1], levels_[level].threads), L0.stream);
            launches++;
            if (e != cudaSuccess || cudaStreamSynchronize(L0.stream) != cudaSuccess) { err = "slow kernel failed"; good = false; }
            if (good && off + W < n) good = setup_centers(L0, level, base.add64(W), err);
        }
        base = base.add64(W);
    }
    if (good && cudaMemcpy(out, d_out, (size_t)n * stride, cudaMemcpyDeviceToHost) != cudaSuccess) { err = "copy of digests failed"; good = false; }
    L0.centers_level = -1;
    lp_ = saved;
    restore();
    return good;
}

bool Engine::bench_resident(const U256 &start, uint32_t n_windows, float &ms, uint64_t &keys, uint64_t &nlaunch, std::string &err) {
    if (!has_pattern_) { err = "no pattern set"; return false; }
    if (lp_.format == VG_FMT_P2TR && !ensure_taproot(~0ULL, err)) return false;      // sustained rate: the long-job table
    if (!compute_stride_points(err)) return false;
    // the lanes of the first device, exactly as a scan drives them: window k of the sequence goes to lane k mod G
    std::vector<Lane *> ls;
    for (auto &L : lanes_) if (L.device == lanes_[0].device) ls.push_back(&L);
    const uint32_t nl = (uint32_t)ls.size();
    CK(cudaSetDevice(ls[0]->device));
    const int fl = (int)levels_.size() - 1;
    const uint64_t W = levels_[fl].window;
    std::vector<vg_scan_params> Ps(nl);
    for (uint32_t j = 0; j < nl; j++) {
        Lane &L = *ls[j];
        uint64_t gl = (uint64_t)shard_index_ * lanes_.size() + j;
        U256 base = start.add(U256(gl).mul64(W));
        if (!setup_centers(L, fl, base, err)) return false;
        Slot &S = L.slot[0];
        S.base = base; S.level = fl; S.limit = W;
        fill_params(L, S, Ps[j]);
        CK(cudaStreamSynchronize(L.stream));
    }
    cudaEvent_t e0, e1; std::vector<cudaEvent_t> done(nl);
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (auto &d : done) CK(cudaEventCreateWithFlags(&d, cudaEventDisableTiming));
    cudaStream_t s0 = ls[0]->stream;
    CK(cudaEventRecord(e0, s0));
    for (uint32_t j = 1; j < nl; j++) CK(cudaStreamWaitEvent(ls[j]->stream, e0, 0));      // no lane starts before e0
    for (uint32_t w = 0; w < n_windows; w++) {
        Lane &L = *ls[w % nl]; Slot &S = L.slot[0];
        CK(cudaMemsetAsync(S.d_counters, 0, kCounterWords * 4, L.stream));
        CK(launch_walk(lp_.format, VG_MODE_MATCH, Ps[w % nl], levels_[fl].blocks, L.stream));
        if (lp_.format == VG_FMT_P2TR) { CK(launch_taptweak(VG_MODE_MATCH, Ps[w % nl], L.stream)); launches++; }
    }
    for (uint32_t j = 1; j < nl; j++) { CK(cudaEventRecord(done[j], ls[j]->stream)); CK(cudaStreamWaitEvent(s0, done[j], 0)); }
    CK(cudaEventRecord(e1, s0));                                                           // after every lane's last window
    CK(cudaEventSynchronize(e1));
    CK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    for (auto &d : done) cudaEventDestroy(d);
    launches += n_windows; nlaunch = n_windows;
    keys = (uint64_t)n_windows * W;
    for (auto *L : ls) L->centers_level = -1;
    return true;
}

}  // namespace vg
