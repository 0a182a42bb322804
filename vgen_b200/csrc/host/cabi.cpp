// extern "C" surface of libvgen_cuda.so (include/vgen_cuda.h) over vg::Engine.
#include "../../../include/vgen_cuda.h"
#include "engine.h"
#include "pattern_lint.h"
#include <string.h>
#include <stdio.h>
#include <new>
#include <thread>
#include <algorithm>

struct vgen_ctx {
    vg::Engine engine;
    std::string last_error;
    std::vector<vg::Hit> hits;     // confirmed hits of the last scan, ascending by key
    std::string text;              // their text records (vg::Hit::text_off / text_len), in hit order
    bool all_text = false;         // every hit has a text record: the records of any hit range are one contiguous piece of `text`
};
static void export_hit(const vg::Hit &h, vgen_hit &o) {
    h.key.to_be(o.key_be);
    memcpy(o.digest, h.digest, 32);
    o.digest_len = h.digest_len; o.device = h.lane;
}
static thread_local std::string g_create_error;

static int fail(vgen_ctx *ctx, int code, const std::string &msg) {
    if (ctx) ctx->last_error = msg; else g_create_error = msg;
    return code;
}
static int copy_out(const std::string &s, char *out, size_t cap) {
    if (!out || cap < s.size() + 1) return VGEN_ERR_BUFFER;
    memcpy(out, s.c_str(), s.size() + 1);
    return (int)s.size();
}

// memcpy of a large blob on several threads (a dense scan's text records are tens of MB; one thread copies ~10 GB/s)
static void parallel_copy(char *dst, const char *src, size_t bytes) {
    const size_t kPiece = 2u << 20;
    const unsigned nt = (unsigned)std::min<size_t>(std::min<size_t>(8, std::max(1u, std::thread::hardware_concurrency())), bytes / kPiece);
    if (nt < 2) { memcpy(dst, src, bytes); return; }
    const size_t per = ((bytes / nt) + 4095) & ~(size_t)4095;
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; t++) {
        const size_t off = (size_t)t * per;
        if (off >= bytes) break;
        th.emplace_back([=] { memcpy(dst + off, src + off, std::min(per, bytes - off)); });
    }
    memcpy(dst, src, std::min(per, bytes));
    for (auto &x : th) x.join();
}

extern "C" {

int vgen_cuda_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}
int vgen_cuda_device_name(int device, char *buf, size_t len) {
    cudaDeviceProp p;
    if (cudaGetDeviceProperties(&p, device) != cudaSuccess) return VGEN_ERR_CUDA;
    return copy_out(p.name, buf, len) < 0 ? VGEN_ERR_BUFFER : VGEN_OK;
}
int vgen_cuda_create(const int *devices, int n_devices, uint32_t hint, vgen_ctx **out) {
    if (!out) return fail(nullptr, VGEN_ERR_INVALID_ARG, "out is NULL");
    *out = nullptr;
    vgen_ctx *ctx = new (std::nothrow) vgen_ctx;
    if (!ctx) return fail(nullptr, VGEN_ERR_INVALID_ARG, "out of memory");
    std::string err;
    try {
        if (!ctx->engine.create(devices, n_devices, hint, err)) { delete ctx; return fail(nullptr, VGEN_ERR_CUDA, err); }
    } catch (const std::exception &e) { delete ctx; return fail(nullptr, VGEN_ERR_CUDA, e.what()); }
    *out = ctx;
    return VGEN_OK;
}
void vgen_cuda_destroy(vgen_ctx *ctx) { delete ctx; }
const char *vgen_cuda_last_error(const vgen_ctx *ctx) { return ctx ? ctx->last_error.c_str() : g_create_error.c_str(); }

int vgen_cuda_set_pattern(vgen_ctx *ctx, int format, const char *pattern, int ci) {
    if (!ctx || !pattern) return VGEN_ERR_INVALID_ARG;
    if (format < 0 || format > 5) return fail(ctx, VGEN_ERR_INVALID_ARG, "invalid address format");
    std::string err;
    try {
        vg::LoweredPattern probe;
        if (!vg::lower_pattern(format, pattern, ci != 0, probe, err)) return fail(ctx, VGEN_ERR_PATTERN, err);
        if (!ctx->engine.set_pattern(format, pattern, ci != 0, err)) return fail(ctx, VGEN_ERR_CUDA, err);
    } catch (const std::exception &e) { return fail(ctx, VGEN_ERR_PATTERN, e.what()); }
    return VGEN_OK;
}
const char *vgen_cuda_matcher_kind(const vgen_ctx *ctx) {
    if (!ctx || !ctx->engine.has_pattern()) return "";
    return ctx->engine.pattern().kind_name.c_str();
}
int vgen_cuda_set_shard(vgen_ctx *ctx, uint32_t index, uint32_t count) {
    if (!ctx) return VGEN_ERR_INVALID_ARG;
    std::string err;
    if (!ctx->engine.set_shard(index, count, err)) return fail(ctx, VGEN_ERR_INVALID_ARG, err);
    return VGEN_OK;
}
int vgen_cuda_set_stream(vgen_ctx *ctx, int lane, void *stream) {
    if (!ctx) return VGEN_ERR_INVALID_ARG;
    std::string err;
    if (!ctx->engine.set_stream(lane, (cudaStream_t)stream, err)) return fail(ctx, VGEN_ERR_CUDA, err);
    return VGEN_OK;
}
int vgen_cuda_scan(vgen_ctx *ctx, const uint8_t start_be[32], const uint8_t *end_be, uint64_t max_hits,
                   const volatile int32_t *stop, vgen_progress_fn progress, void *user,
                   vgen_hit *hits_out, uint64_t hits_cap, uint64_t *n_hits, uint64_t *keys_scanned) {
    if (!ctx || !start_be || !n_hits || !keys_scanned || (!hits_out && hits_cap)) return VGEN_ERR_INVALID_ARG;
    *n_hits = 0; *keys_scanned = 0;
    std::string err;
    try {
        vg::U256 start = vg::U256::from_be(start_be), end;
        if (end_be) { end = vg::U256::from_be(end_be); if (end < start) return fail(ctx, VGEN_ERR_RANGE, "range end is below range start"); }
        std::vector<vg::Hit> &hits = ctx->hits;
        hits.clear();
        vg::Engine::Progress cb;
        if (progress) cb = [&](uint64_t k) { progress(k, user); };
        if (!ctx->engine.scan(start, end_be ? &end : nullptr, max_hits, stop, cb, hits, ctx->text, *keys_scanned, err)) return fail(ctx, VGEN_ERR_CUDA, err);
        ctx->all_text = true;
        for (const vg::Hit &h : hits) if (h.text_len == 0) { ctx->all_text = false; break; }
        uint64_t n = std::min<uint64_t>(hits.size(), hits_cap);
        for (uint64_t i = 0; i < n; i++) export_hit(hits[i], hits_out[i]);
        *n_hits = n;
    } catch (const std::exception &e) { return fail(ctx, VGEN_ERR_CUDA, e.what()); }
    return VGEN_OK;
}
uint64_t vgen_cuda_scan_total_hits(const vgen_ctx *ctx) { return ctx ? ctx->hits.size() : 0; }
uint64_t vgen_cuda_scan_fetch_hits(const vgen_ctx *ctx, uint64_t offset, vgen_hit *out, uint64_t cap) {
    if (!ctx || !out || offset >= ctx->hits.size()) return 0;
    uint64_t n = std::min<uint64_t>(ctx->hits.size() - offset, cap);
    for (uint64_t i = 0; i < n; i++) export_hit(ctx->hits[offset + i], out[i]);
    return n;
}
uint64_t vgen_cuda_scan_fetch_records(const vgen_ctx *ctx, uint64_t offset, uint64_t cap, uint8_t *keys32, char *addresses, char *wifs, size_t stride) {
    if (!ctx || offset >= ctx->hits.size() || stride < 96) return 0;
    const int fmt = ctx->engine.pattern().format;
    uint64_t n = std::min<uint64_t>(ctx->hits.size() - offset, cap);
    for (uint64_t i = 0; i < n; i++) {
        const vg::Hit &h = ctx->hits[offset + i];
        uint8_t kb[32]; h.key.to_be(kb);
        if (keys32) memcpy(keys32 + 32 * i, kb, 32);
        if (addresses) { std::string a = vg::format_address(fmt, h.digest); strncpy(addresses + stride * i, a.c_str(), stride - 1); addresses[stride * i + stride - 1] = 0; }
        if (wifs) { std::string w = vg::format_wif(fmt, kb); strncpy(wifs + stride * i, w.c_str(), stride - 1); wifs[stride * i + stride - 1] = 0; }
    }
    return n;
}
// "<key hex> <address> <wif>\n" per hit, formatted on all host threads (the reference builds these per match with rayon,
// src/gpu.rs:1030-1093)
int64_t vgen_cuda_scan_fetch_text(const vgen_ctx *ctx, uint64_t offset, uint64_t max_hits, char *buf, size_t cap, uint64_t *n_written) {
    if (!ctx || !buf || !n_written) return VGEN_ERR_INVALID_ARG;
    *n_written = 0;
    if (offset >= ctx->hits.size()) { if (cap) buf[0] = 0; return 0; }
    const int fmt = ctx->engine.pattern().format;
    const size_t kRec = 64 + 1 + 96 + 1 + 64 + 1;                 // upper bound of one line
    uint64_t n = std::min<uint64_t>(ctx->hits.size() - offset, max_hits);
    if (cap < 1) return VGEN_ERR_BUFFER;
    if (n == 0) return VGEN_ERR_BUFFER;
    if (ctx->all_text) {
        // the scan's confirm jobs built every record: the records of hits [offset, offset + n) are one piece of the text blob
        const vg::Hit *H = ctx->hits.data() + offset;
        const uint64_t t0 = H[0].text_off;
        uint64_t lo = 0, hi = n;                                   // largest count whose bytes fit in cap - 1
        while (lo < hi) { uint64_t mid = (lo + hi + 1) / 2; if (H[mid - 1].text_off + H[mid - 1].text_len - t0 <= cap - 1) lo = mid; else hi = mid - 1; }
        if (lo == 0) return VGEN_ERR_BUFFER;
        const size_t bytes = (size_t)(H[lo - 1].text_off + H[lo - 1].text_len - t0);
        parallel_copy(buf, ctx->text.data() + t0, bytes);
        buf[bytes] = 0;
        *n_written = lo;
        return (int64_t)bytes;
    }
    n = std::min<uint64_t>(n, (cap - 1) / kRec);
    if (n == 0) return VGEN_ERR_BUFFER;
    std::vector<std::string> lines(n);
    unsigned nt = std::max(1u, std::min<unsigned>(std::thread::hardware_concurrency(), (unsigned)(n / 256 + 1)));
    auto work = [&](unsigned t) {
        for (uint64_t i = t; i < n; i += nt) {
            const vg::Hit &h = ctx->hits[offset + i];
            if (h.text_len) { lines[i].assign(ctx->text, (size_t)h.text_off, h.text_len); continue; }      // built by the scan's confirm job
            uint8_t kb[32]; h.key.to_be(kb);
            lines[i] = vg::h_hex(kb, 32) + " " + vg::format_address(fmt, h.digest) + " " + vg::format_wif(fmt, kb) + "\n";
        }
    };
    std::vector<std::thread> th;
    for (unsigned t = 1; t < nt; t++) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    size_t pos = 0;
    for (uint64_t i = 0; i < n; i++) { memcpy(buf + pos, lines[i].data(), lines[i].size()); pos += lines[i].size(); }
    buf[pos] = 0;
    *n_written = n;
    return (int64_t)pos;
}
int vgen_cuda_scan_text(const vgen_ctx *ctx, const char **text, uint64_t *len) {
    if (!ctx || !text || !len || !ctx->all_text) return 0;
    // hits beyond max_hits were cut off the hit list; their records are still at the end of the blob: stop at the last kept hit
    const uint64_t n = ctx->hits.empty() ? 0 : ctx->hits.back().text_off + ctx->hits.back().text_len;
    *text = ctx->text.c_str(); *len = n;
    return 1;
}
int vgen_cuda_digest_batch(vgen_ctx *ctx, int format, const uint8_t start_be[32], uint64_t n, uint8_t *out, uint32_t stride) {
    if (!ctx || !start_be || (!out && n)) return VGEN_ERR_INVALID_ARG;
    std::string err;
    try {
        if (!ctx->engine.digest_batch(format, vg::U256::from_be(start_be), n, out, stride, err)) return fail(ctx, VGEN_ERR_CUDA, err);
    } catch (const std::exception &e) { return fail(ctx, VGEN_ERR_CUDA, e.what()); }
    return VGEN_OK;
}
int vgen_cuda_hash160_batch(vgen_ctx *ctx, const uint8_t start_be[32], uint64_t n, uint8_t *out20) {
    return vgen_cuda_digest_batch(ctx, VGEN_P2PKH, start_be, n, out20, 20);
}
int vgen_cuda_bench_resident(vgen_ctx *ctx, const uint8_t start_be[32], uint32_t n_windows, float *ms, uint64_t *keys, uint64_t *launches) {
    if (!ctx || !start_be || !ms || !keys || !launches) return VGEN_ERR_INVALID_ARG;
    std::string err;
    if (!ctx->engine.bench_resident(vg::U256::from_be(start_be), n_windows, *ms, *keys, *launches, err)) return fail(ctx, VGEN_ERR_CUDA, err);
    return VGEN_OK;
}
uint64_t vgen_cuda_launch_count(const vgen_ctx *ctx) { return ctx ? ctx->engine.launches.load() : 0; }
void vgen_cuda_io_bytes(const vgen_ctx *ctx, uint64_t *h2d, uint64_t *d2h) {
    if (h2d) *h2d = ctx ? ctx->engine.h2d_bytes.load() : 0;
    if (d2h) *d2h = ctx ? ctx->engine.d2h_bytes.load() : 0;
}
void vgen_cuda_stats(const vgen_ctx *ctx, uint64_t out[8]) {
    if (!out) return;
    memset(out, 0, 8 * sizeof(uint64_t));
    if (!ctx) return;
    const vg::Engine &e = ctx->engine;
    out[0] = (uint64_t)e.n_devices(); out[1] = (uint64_t)e.n_failed_devices(); out[2] = e.requeued_windows.load();
    out[3] = e.overflow_rescans.load(); out[4] = e.hit_cap(); out[5] = e.pool_threads();
    auto g = const_cast<vg::Engine &>(e).check_guards();
    out[6] = g.first; out[7] = g.second;
}
int vgen_cuda_lane_count(const vgen_ctx *ctx) { return ctx ? ctx->engine.n_lanes() : 0; }
int vgen_cuda_tap_bits(const vgen_ctx *ctx) { return ctx ? (int)ctx->engine.tap_bits() : 0; }
uint64_t vgen_cuda_window_keys(const vgen_ctx *ctx) { return ctx ? ctx->engine.levels().back().window : 0; }

int vgen_format_address(int format, const uint8_t *digest, char *out, size_t cap) {
    if (!digest || format < 0 || format > 5) return VGEN_ERR_INVALID_ARG;
    return copy_out(vg::format_address(format, digest), out, cap);
}
int vgen_format_wif(int format, const uint8_t key_be[32], char *out, size_t cap) {
    if (!key_be || format < 0 || format > 5) return VGEN_ERR_INVALID_ARG;
    return copy_out(vg::format_wif(format, key_be), out, cap);
}
int vgen_pattern_check(const char *pattern, int ci, char *errbuf, size_t cap) {
    if (!pattern) return VGEN_ERR_INVALID_ARG;
    vg::Dfa d; vg::PatternInfo info; std::string err;
    try {
        if (vg::compile_pattern(pattern, ci != 0, d, info, err)) return VGEN_OK;
    } catch (const std::exception &e) { err = e.what(); }
    if (errbuf && cap) { strncpy(errbuf, err.c_str(), cap - 1); errbuf[cap - 1] = 0; }
    return VGEN_ERR_PATTERN;
}
int vgen_pattern_is_match(const char *pattern, int ci, const char *text) {
    if (!pattern || !text) return VGEN_ERR_INVALID_ARG;
    vg::Dfa d; vg::PatternInfo info; std::string err;
    try {
        if (!vg::compile_pattern(pattern, ci != 0, d, info, err)) return VGEN_ERR_PATTERN;
    } catch (const std::exception &) { return VGEN_ERR_PATTERN; }
    return d.match(text, strlen(text)) ? 1 : 0;
}

int64_t vgen_plan_windows(const uint8_t start_be[32], const uint8_t end_be[32], int sm_count, uint32_t keys_per_launch_hint,
                          uint32_t global_lanes, uint32_t lane, int can_end_early, uint8_t *first_keys_be, uint64_t *n_keys, uint64_t cap) {
    if (!start_be || !end_be || sm_count <= 0 || global_lanes == 0 || lane >= global_lanes) return VGEN_ERR_INVALID_ARG;
    vg::U256 start = vg::U256::from_be(start_be), end = vg::U256::from_be(end_be);
    if (end < start) return VGEN_ERR_RANGE;
    if (start.is_zero()) start = vg::U256(1);
    if (vg::secp_order_minus1() < end) end = vg::secp_order_minus1();
    if (end < start) return 0;
    const vg::U256 N = end.sub(start).add64(1);
    std::vector<vg::Level> levels = vg::compute_levels(sm_count, keys_per_launch_hint, 512, VG_BLOCK, VG_CTAS_PER_SM);
    vg::WindowPlan plan(levels, global_lanes, N, can_end_early != 0);
    vg::LaneCursor c = plan.begin();
    int64_t n = 0;
    vg::U256 off; uint64_t limit;
    while (plan.window(lane, c, off, limit)) {
        if ((uint64_t)n < cap) {
            if (first_keys_be) start.add(off).to_be(first_keys_be + 32 * n);
            if (n_keys) n_keys[n] = limit;
        }
        n++;
        plan.advance(c);
        if (n > (1 << 24)) break;
    }
    return n;
}
int vgen_pattern_dfa(int format, const char *pattern, int ci, uint16_t *table, size_t table_cap, uint8_t *accept, size_t accept_cap,
                     uint8_t sym_class[64], uint32_t info[4]) {
    if (!pattern || !info || format < 0 || format > 5) return VGEN_ERR_INVALID_ARG;
    vg::LoweredPattern lp; std::string err;
    try {
        if (!vg::lower_pattern(format, pattern, ci != 0, lp, err)) return VGEN_ERR_PATTERN;
    } catch (const std::exception &) { return VGEN_ERR_PATTERN; }
    info[0] = lp.dev.dfa.n_states; info[1] = lp.dev.dfa.n_classes; info[2] = lp.dev.dfa.start; info[3] = lp.dev.dfa.flags;
    if (sym_class) memcpy(sym_class, lp.dev.dfa.sym_class, 64);
    // the exact DFA only (an Ethereum case-folded prefilter DFA may be appended behind it, lowering.cpp)
    const size_t tn = (size_t)lp.dfa.n_states * lp.dfa.n_classes, an = lp.dfa.n_states;
    if (!table || !accept || table_cap < tn || accept_cap < an) return VGEN_ERR_BUFFER;
    memcpy(table, lp.dfa.table.data(), tn * 2);
    memcpy(accept, lp.dfa.accept_end.data(), an);
    return VGEN_OK;
}
int vgen_pattern_lowering(int format, const char *pattern, int ci, char *out, size_t cap) {
    if (!pattern || format < 0 || format > 5) return VGEN_ERR_INVALID_ARG;
    vg::LoweredPattern lp; std::string err;
    try {
        if (!vg::lower_pattern(format, pattern, ci != 0, lp, err)) return VGEN_ERR_PATTERN;
    } catch (const std::exception &) { return VGEN_ERR_PATTERN; }
    std::string s = lp.kind_name;
    char buf[64];
    if (lp.dev.kind == VG_MATCH_INTERVALS)
        for (int i = 0; i < lp.dev.n_intervals; i++) {
            snprintf(buf, sizeof buf, ";%08x%08x-%08x%08x", lp.dev.lo_hi[i][0], lp.dev.lo_hi[i][1], lp.dev.lo_hi[i][2], lp.dev.lo_hi[i][3]);
            s += buf;
        }
    if (lp.dev.kind == VG_MATCH_EXACT) {
        s += ";exact=";
        for (int i = 0; i < lp.dev.exact_words * 4; i++) { snprintf(buf, sizeof buf, "%02x", (lp.dev.exact[i / 4] >> (8 * (i % 4))) & 0xff); s += buf; }
    }
    if (lp.host_regex) s += ";confirm=host-nfa";      // DFA too large for the device: superset on the device, exact NFA match on the host
    if (lp.dev.confirm_dfa) s += ";confirm=dfa";       // the device runs the DFA on the keys that pass the prefilter
    if (lp.dev.kind == VG_MATCH_DFA && lp.dev.dfa.chk_mask != 0) {      // literal-suffix test ahead of the DFA (scan_types.h)
        snprintf(buf, sizeof buf, ";suffix=%08x:%08x", lp.dev.dfa.chk_mask, lp.dev.dfa.chk_value);
        s += buf;
    }
    if (lp.dev.kind == VG_MATCH_DFA && lp.dev.dfa.b58_mod != 0) {      // Base58 literal suffix: payload mod 58^k in a small set
        snprintf(buf, sizeof buf, ";residue=%x", lp.dev.dfa.b58_mod);
        s += buf;
        for (uint32_t i = 0; i < lp.dev.dfa.b58_nvals; i++) { snprintf(buf, sizeof buf, ":%x", lp.dev.dfa.b58_vals[i]); s += buf; }
    }
    return copy_out(s, out, cap);
}

// ---- CLI support: lint, difficulty, provider, key parsing (host only) ----
int vgen_pattern_validate_charset(int format, const char *pattern, int ci, char *out, size_t cap) {
    if (!pattern || format < 0 || format > 5) return VGEN_ERR_INVALID_ARG;
    return copy_out(vg::validate_charset(pattern, ci != 0, format), out, cap);
}
uint64_t vgen_pattern_estimate_difficulty(int format, const char *pattern, int ci) {
    if (!pattern || format < 0 || format > 5) return 0;
    return vg::estimate_difficulty(pattern, ci != 0, format);
}
int vgen_provider_resolve(const char *ref, char *address, size_t cap, int *format, uint8_t start_be[32], uint8_t end_be[32],
                          int *has_range, char *errbuf, size_t err_cap) {
    if (!ref || !address || !format || !start_be || !end_be || !has_range) return VGEN_ERR_INVALID_ARG;
    vg::ProviderResult r; std::string err;
    int rc = vg::provider_resolve(ref, r, err);
    if (rc < 0) { if (errbuf && err_cap) { strncpy(errbuf, err.c_str(), err_cap - 1); errbuf[err_cap - 1] = 0; } return VGEN_ERR_PATTERN; }
    if (rc == 0) return 0;
    if (copy_out(r.address, address, cap) < 0) return VGEN_ERR_BUFFER;
    *format = r.format; *has_range = r.has_range ? 1 : 0;
    r.start.to_be(start_be); r.end.to_be(end_be);
    return 1;
}
int vgen_provider_pattern(const char *address, int prefix_len, char *out, size_t cap) {
    if (!address) return VGEN_ERR_INVALID_ARG;
    vg::ProviderResult r; r.address = address;
    return copy_out(prefix_len > 0 ? vg::provider_build_pattern(r, (size_t)prefix_len) : vg::provider_build_exact_pattern(r), out, cap);
}
// WIF (mainnet, compressed or not) or 64 hex digits (optional 0x) -> key; the `verify` command's parser (src/lib.rs:387-401)
int vgen_parse_key(const char *text, uint8_t key_be[32], int *is_wif, int *compressed) {
    if (!text || !key_be) return VGEN_ERR_INVALID_ARG;
    std::string s = text;
    std::vector<uint8_t> payload;
    if (vg::h_base58check_decode(s, payload) && (payload.size() == 33 || (payload.size() == 34 && payload[33] == 0x01)) && payload[0] == 0x80) {
        memcpy(key_be, payload.data() + 1, 32);
        if (is_wif) *is_wif = 1;
        if (compressed) *compressed = payload.size() == 34;
    } else {
        while (s.size() >= 2 && s[0] == '0' && s[1] == 'x') s = s.substr(2);           // trim_start_matches("0x")
        if (s.size() != 64) return VGEN_ERR_INVALID_ARG;
        for (int i = 0; i < 32; i++) {
            int v = 0;
            for (int j = 0; j < 2; j++) {
                int c = s[2 * i + j], d;
                if (c >= '0' && c <= '9') d = c - '0'; else if (c >= 'a' && c <= 'f') d = c - 'a' + 10; else if (c >= 'A' && c <= 'F') d = c - 'A' + 10; else return VGEN_ERR_INVALID_ARG;
                v = v * 16 + d;
            }
            key_be[i] = (uint8_t)v;
        }
        if (is_wif) *is_wif = 0;
        if (compressed) *compressed = 1;
    }
    vg::U256 k = vg::U256::from_be(key_be);
    if (k.is_zero() || vg::secp_order_minus1() < k) return VGEN_ERR_RANGE;             // SecretKey::from_slice rejects 0 and >= n
    return VGEN_OK;
}

}  // extern "C"
