// Window schedule of a scan: pure host arithmetic (no CUDA), shared by the engine and by the CPU tests of the
// multi-rank path. With G global lanes (ranks x GPUs per rank) the range [0, N) is cut into a fixed sequence of
// windows: one window per lane at each ramp level first_level..final-1, then full-size windows handed out
// round-robin (lane g takes full windows g, g+G, g+2G, ...). Every lane therefore owns disjoint windows and walks
// the full-size part with the common stride G*W; no lane ever needs another lane's data.
// Reference counterpart: the single sequential window loop of scan_gpu_with_runner (src/gpu.rs:973-1028).
#pragma once
#include <vector>
#include "host_util.h"

namespace vg {

struct Level { uint32_t blocks; uint32_t threads; uint64_t window; };   // launch geometry of one ramp level

// levels for a device with `sm_count` SMs: optional small ramp levels, then the full-size level
inline std::vector<Level> compute_levels(int sm_count, uint32_t keys_per_launch_hint, uint32_t half, uint32_t block, uint32_t ctas_per_sm) {
    uint32_t full_blocks = (uint32_t)sm_count * ctas_per_sm;       // one full wave of resident CTAs
    if (keys_per_launch_hint) {
        uint32_t b = keys_per_launch_hint / (block * 2 * half);
        full_blocks = b < 1 ? 1 : (b > 8u * sm_count ? 8u * sm_count : b);
    }
    std::vector<Level> lv;
    for (uint32_t b : {1u, 8u, 64u}) if (b * 4 <= full_blocks) lv.push_back({b, b * block, (uint64_t)b * block * 2 * half});
    lv.push_back({full_blocks, full_blocks * block, (uint64_t)full_blocks * block * 2 * half});
    return lv;
}

struct LaneCursor { int level = 0; uint64_t k = 0; };

class WindowPlan {
public:
    // can_end_early: a hit limit on a matcher that may be dense (the ramp pays off); otherwise the plan starts at the
    // smallest level that covers N in one round
    WindowPlan(const std::vector<Level> &levels, uint32_t global_lanes, const U256 &N, bool can_end_early)
        : levels_(levels), G_(global_lanes), N_(N), base_(levels.size()) {
        const int fl = (int)levels_.size() - 1;
        first_ = 0;
        if (!can_end_early) {
            first_ = fl;
            for (int lv = 0; lv < fl; lv++) if (!(U256(levels_[lv].window).mul64(G_) < N_)) { first_ = lv; break; }
        }
        for (int lv = first_ + 1; lv <= fl; lv++) base_[lv] = base_[lv - 1].add(U256(levels_[lv - 1].window).mul64(G_));
    }
    int first_level() const { return first_; }
    int final_level() const { return (int)levels_.size() - 1; }
    LaneCursor begin() const { LaneCursor c; c.level = first_; return c; }
    // window the cursor points at for global lane gl: offset from the range start, keys inside the range. false = lane done.
    bool window(uint32_t gl, const LaneCursor &c, U256 &off, uint64_t &limit) const {
        const int fl = final_level();
        U256 idx = c.level < fl ? U256((uint64_t)gl) : U256(c.k).mul64(G_).add64(gl);
        off = base_[c.level].add(idx.mul64(levels_[c.level].window));
        if (!(off < N_)) return false;
        U256 remain = N_.sub(off);
        limit = (remain.fits64() && remain.w[0] < levels_[c.level].window) ? remain.w[0] : levels_[c.level].window;
        return true;
    }
    void advance(LaneCursor &c) const { if (c.level < final_level()) c.level++; else c.k++; }
    // scalar distance between consecutive windows of one lane at `level` (what the walk adds to every centre)
    U256 stride(int level) const { return U256(levels_[level].window).mul64(G_); }
private:
    std::vector<Level> levels_;
    uint32_t G_;
    U256 N_;
    std::vector<U256> base_;
    int first_;
};

}  // namespace vg
