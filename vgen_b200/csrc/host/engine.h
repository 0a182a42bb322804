// Host driver of the scan path: owns per-GPU lanes (stream, centres, scratch, hit buffers), runs the
// window loop, merges and confirms hits. Replaces GpuRunner + scan_gpu_with_runner
// (src/gpu.rs:134-660, 920-1125). One scan at a time per engine.
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include <functional>
#include "lowering.h"
#include "plan.h"
#include "../kernels.h"

namespace vg {

// text: "<key hex> <address> <wif>\n", filled by the per-window confirm job when the window is sparse (else empty: formatted on demand)
struct Hit { U256 key; uint8_t digest[32]; uint32_t digest_len; uint32_t lane; std::string text; };

struct Slot {                 // one of the two windows a lane keeps in flight
    vg_hit_rec *d_hits = nullptr;
    uint32_t *d_counters = nullptr;    // [0] hit_count, [1] degen_count
    uint32_t *d_degen = nullptr;
    uint32_t *d_base8 = nullptr;       // window base scalar for the device-chained slow pass (P2TR)
    uint32_t *h_counters = nullptr;    // pinned
    cudaEvent_t done = nullptr;
    bool busy = false;
    U256 base;                         // first scalar of the window
    uint64_t limit = 0;                // keys of the window inside the range
    int level = 0;
    uint64_t seq = 0;                  // global window sequence number
};

struct Lane {
    int device = 0;
    cudaStream_t stream = nullptr, copy_stream = nullptr;
    bool own_stream = true;
    int sm_count = 0;
    uint32_t *d_pow2 = nullptr, *d_gtab = nullptr, *d_centers = nullptr, *d_scratch = nullptr, *d_base8 = nullptr;
    bool owns_tap_table = false;                 // the lanes of one GPU share the table
    uint32_t *d_points = nullptr, *d_tap_table = nullptr;   // P2TR only, allocated on first use (taproot.cuh)
    uint32_t *d_stride_pts = nullptr;          // [n_levels][16] scratch for stride points
    uint16_t *d_dfa_table = nullptr; uint8_t *d_dfa_accept = nullptr;
    vg_hit_rec *h_hits = nullptr; size_t h_hits_cap = 0;   // pinned staging
    Slot slot[2];
    std::vector<std::array<uint32_t, 16>> stride_pt;       // per level, for the current lane count
    int centers_level = -1;                    // level the centres are currently set up for (-1: none)
    std::string name;
};

class Engine {
public:
    Engine() {}
    ~Engine();
    bool create(const int *devices, int n_devices, uint32_t keys_per_launch_hint, std::string &err);
    bool set_pattern(int format, const std::string &pattern, bool ci, std::string &err);
    bool set_shard(uint32_t index, uint32_t count, std::string &err);
    bool set_stream(int lane, cudaStream_t s, std::string &err);
    typedef std::function<void(uint64_t)> Progress;
    // scans [start, end] (end == nullptr: to n-1); hits sorted ascending, truncated to max_hits (0 = unlimited)
    bool scan(const U256 &start, const U256 *end, uint64_t max_hits, const volatile int32_t *stop, const Progress &progress,
              std::vector<Hit> &hits, uint64_t &keys_scanned, std::string &err);
    // digests of n sequential keys from start (stride bytes each, zero-filled for invalid scalars)
    bool digest_batch(int format, const U256 &start, uint64_t n, uint8_t *out, uint32_t stride, std::string &err);
    // device-resident timing of n_windows full-size windows on lane 0 (bench `value` and roofline)
    bool bench_resident(const U256 &start, uint32_t n_windows, float &ms, uint64_t &keys, uint64_t &launches, std::string &err);
    const LoweredPattern &pattern() const { return lp_; }
    bool has_pattern() const { return has_pattern_; }
    static int lanes_per_device();      // lanes (streams) the engine opens on every GPU (default 2, VGEN_LANES_PER_GPU)
    int n_lanes() const { return (int)lanes_.size(); }
    uint32_t tap_bits() const { return tap_bits_; }
    const Lane &lane(int i) const { return lanes_[i]; }
    uint32_t half() const { return half_; }
    const std::vector<Level> &levels() const { return levels_; }
    uint64_t launches = 0;        // kernel launches issued so far (walk + slow + set-up)
    uint64_t h2d_bytes = 0, d2h_bytes = 0;   // bytes copied host->device / device->host by scans so far
private:
    bool setup_centers(Lane &L, int level, const U256 &window_base, std::string &err);
    bool compute_stride_points(std::string &err);
    bool launch_window(Lane &L, int s, int fmt, int mode, uint8_t *dump, uint32_t dump_stride, std::string &err);
    void fill_params(Lane &L, Slot &S, vg_scan_params &P);
    bool upload_dfa(std::string &err);
    bool ensure_taproot(uint64_t keys_hint, std::string &err);
    uint32_t tap_bits_ = 0;                      // window width of the P2TR table currently built (0 = none)
    void destroy();
    std::vector<Lane> lanes_;
    std::vector<Level> levels_;
    uint32_t half_ = 512;
    uint32_t shard_index_ = 0, shard_count_ = 1;
    uint32_t stride_lanes_ = 0;    // lane count the stride points were computed for
    LoweredPattern lp_;
    bool has_pattern_ = false;
    bool p2tr_internal_ = false;   // digest_batch only: dump the untweaked internal keys instead of running the tweak stage
};

}  // namespace vg
