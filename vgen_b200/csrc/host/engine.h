// Host driver of the scan path: owns the per-GPU tables and the per-lane buffers (stream, centres, scratch, hit buffers),
// runs the window loop with ONE HOST THREAD PER DEVICE, confirms hits on a worker pool and merges them, in window order,
// on the calling thread. Replaces GpuRunner + scan_gpu_with_runner (src/gpu.rs:134-660, 920-1125). One scan at a time per
// engine.
#pragma once
#include <cuda_runtime.h>
#include <array>
#include <atomic>
#include <memory>
#include <condition_variable>
#include <deque>
#include <functional>
#include <future>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "lowering.h"
#include "plan.h"
#include "../kernels.h"

namespace vg {

// A confirmed hit. Its text record "<key hex> <address> <wif>\n" is built by the per-window confirm job when the window is sparse
// and lives in the scan's text blob at [text_off, text_off + text_len) (text_len == 0: formatted on demand). Plain data, so that
// merging a window is two bulk appends (records, text) whatever the number of hits.
struct Hit { U256 key; uint8_t digest[32]; uint32_t digest_len; uint32_t lane; uint64_t text_off; uint32_t text_len; };
// what one confirm job hands back: hits in ascending key order, text offsets relative to `text`
struct HitBatch { std::vector<Hit> hits; std::string text; };

// One window of the fixed window sequence of a scan (plan.h). `seq` orders the windows of this context by key, `logical` is
// the plan lane the window belongs to (normally run by the physical lane of the same index; by any surviving lane once its
// device has failed).
struct WindowTask { uint64_t seq = 0; uint32_t logical = 0; U256 base; uint64_t limit = 0; int level = 0; };

struct Slot {                 // one of the two windows a lane keeps in flight
    vg_hit_rec *d_hits = nullptr;
    uint32_t *d_counters = nullptr;    // [0] hit_count, [1] degen_count
    uint32_t *d_degen = nullptr;
    uint32_t *d_base8 = nullptr;       // window base scalar for the slow pass over degenerate batches
    uint32_t *h_counters = nullptr;    // pinned
    vg_hit_rec *h_hits = nullptr;      // pinned staging of this slot (kStageRecords records)
    cudaEvent_t done = nullptr;
    bool busy = false;
    WindowTask task;
};

struct Lane {
    int dev = 0;                                 // index into Engine::devices_
    int device = 0;                              // CUDA device ordinal
    cudaStream_t stream = nullptr, copy_stream = nullptr;
    bool own_stream = true;
    uint32_t *d_centers = nullptr, *d_scratch = nullptr, *d_base8 = nullptr;
    uint32_t *d_points = nullptr;                // P2TR only, allocated on first use (taproot.cuh)
    uint32_t *d_stride_pts = nullptr;            // [16] scratch for one stride point
    Slot slot[2];
    std::vector<std::array<uint32_t, 16>> stride_pt;       // per level, for the current global lane count
    int centers_level = -1;                      // level the centres are set up for (-1: none)
    U256 centers_base;                           // window base they stand at (valid when centers_level >= 0)
};

// What the lanes of one GPU share: the generator tables, the P2TR fixed-base table, the DFA of the current pattern.
struct Device {
    int device = 0, sm_count = 0;
    std::string name;
    uint32_t *d_pow2 = nullptr, *d_gtab = nullptr, *d_tap_table = nullptr;
    uint16_t *d_dfa_table = nullptr; uint8_t *d_dfa_accept = nullptr;
    size_t dfa_table_cap = 0, dfa_accept_cap = 0;   // bytes allocated (buffers are reused by later patterns that fit)
    std::vector<int> lanes;                      // indices into Engine::lanes_
    bool failed = false;                         // a CUDA call failed on it (or a fault was injected): its windows go to the other devices
    std::string failure;
};

// Fixed pool of host threads for the per-window confirm jobs (re-encode + full DFA + text of every device hit).
class WorkerPool {
public:
    explicit WorkerPool(unsigned n);
    ~WorkerPool();
    std::future<HitBatch> submit(std::function<HitBatch()> fn);
    unsigned size() const { return (unsigned)threads_.size(); }
private:
    void run();
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<std::packaged_task<HitBatch()>> q_;
    std::vector<std::thread> threads_;
    bool stop_ = false;
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
    // `text`: the text records of the hits that have one (Hit::text_off / text_len), in hit order
    bool scan(const U256 &start, const U256 *end, uint64_t max_hits, const volatile int32_t *stop, const Progress &progress,
              std::vector<Hit> &hits, std::string &text, uint64_t &keys_scanned, std::string &err);
    // digests of n sequential keys from start (stride bytes each, zero-filled for invalid scalars)
    bool digest_batch(int format, const U256 &start, uint64_t n, uint8_t *out, uint32_t stride, std::string &err);
    // device-resident timing of n_windows full-size windows per device, all devices of the context at once (bench `value`
    // and roofline): ms = the slowest device's CUDA-event time, keys / launches = totals over the devices
    bool bench_resident(const U256 &start, uint32_t n_windows, float &ms, uint64_t &keys, uint64_t &launches, std::string &err);
    const LoweredPattern &pattern() const { return lp_; }
    bool has_pattern() const { return has_pattern_; }
    static int lanes_per_device();      // lanes (streams) the engine opens on every GPU (default 2, VGEN_LANES_PER_GPU)
    int n_lanes() const { return (int)lanes_.size(); }
    int n_devices() const { return (int)devices_.size(); }
    int n_failed_devices() const { int n = 0; for (auto &d : devices_) n += d.failed; return n; }
    uint32_t tap_bits() const { return tap_bits_; }
    const Lane &lane(int i) const { return lanes_[i]; }
    const Device &device(int i) const { return devices_[i]; }
    uint32_t half() const { return half_; }
    uint32_t hit_cap() const { return hit_cap_; }
    unsigned pool_threads() const { return pool_ ? pool_->size() : 0; }
    const std::vector<Level> &levels() const { return levels_; }
    std::atomic<uint64_t> launches{0};        // kernel launches issued so far (walk + slow + set-up)
    std::atomic<uint64_t> h2d_bytes{0}, d2h_bytes{0};   // bytes copied host->device / device->host by scans so far
    std::atomic<uint64_t> overflow_rescans{0};          // windows whose hits exceeded the hit buffer and were re-run in pieces
    std::atomic<uint64_t> requeued_windows{0};          // windows taken over from a failed device
    // VGEN_GUARD=1 (tools/check_build.sh): every device buffer gets a 256-byte canary behind it; returns (regions, corrupted)
    std::pair<uint64_t, uint64_t> check_guards();
private:
    struct ScanRun;
    struct Harvest;
    bool setup_centers(Lane &L, int level, const U256 &window_base, std::string &err);
    bool compute_stride_points(std::string &err);
    bool launch_window(Lane &L, Slot &S, int fmt, int mode, uint8_t *dump, uint32_t dump_stride, const uint32_t *stride16, std::string &err);
    void fill_params(Lane &L, Slot &S, const uint32_t *stride16, vg_scan_params &P);
    bool harvest(Lane &L, Slot &S, int fmt, ScanRun &R, Harvest &out, std::string &err);
    bool fetch_records(Lane &L, Slot &S, uint32_t n_hits, const U256 &base, uint64_t limit, Harvest &out, std::string &err);
    void device_worker(ScanRun &R, int dev_index);
    bool upload_dfa(std::string &err);
    cudaError_t dmalloc(void **p, size_t bytes, int device);       // cudaMalloc (+ canary under VGEN_GUARD)
    struct Guard { int device; uint8_t *at; };
    std::vector<Guard> guards_;
    std::mutex guards_mu_;
    bool guard_ = false;
    bool ensure_taproot(uint64_t keys_hint, std::string &err);
    uint32_t tap_bits_ = 0;                      // window width of the P2TR table currently built (0 = none)
    void destroy();
    std::vector<Device> devices_;
    std::vector<Lane> lanes_;
    std::vector<Level> levels_;
    std::vector<std::array<uint32_t, 16>> stride1_;   // per level: W*G, the stride of a single lane walking alone (digest_batch)
    uint32_t half_ = 512;
    uint32_t hit_cap_ = 0;         // records per slot hit buffer; a window reporting more is re-run in pieces (harvest)
    int rescan_level_ = 0;         // largest level whose window fits the hit buffer
    uint32_t shard_index_ = 0, shard_count_ = 1;
    uint32_t stride_lanes_ = 0;    // lane count the stride points were computed for
    LoweredPattern lp_;
    bool has_pattern_ = false;
    bool dfa_uploaded_ = false;    // the devices hold lp_'s DFA
    bool p2tr_internal_ = false;   // digest_batch only: dump the untweaked internal keys instead of running the tweak stage
    std::unique_ptr<WorkerPool> pool_;
};

}  // namespace vg
