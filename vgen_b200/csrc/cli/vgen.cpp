// vgen — command-line front end of the B200 scan path, mirroring the reference's CLI surface
// (src/lib.rs:44-211 arguments, :273-558 command handlers, :664-976 run_search + writers, :979-1036 list-gpus,
// :1038-1086 formatting helpers). Uses only the C ABI in include/vgen_cuda.h.
//
// Differences from the reference, all forced by BASELINE.json's north_star ("no CPU fallback"):
//   * --no-gpu is a hard error (the reference's rayon CPU scanner is not part of this product);
//   * Ethereum and P2TR scan on the GPU (the reference falls back to the CPU for Ethereum, src/lib.rs:316-319);
//   * no TUI: --tui / --no-tui are accepted and ignored; progress goes to stderr (plain, or JSON lines with --progress-json);
//   * range scans can be checkpointed and resumed (--checkpoint FILE), and spread over several GPUs (--devices 0,1,...).
#include "../../../include/vgen_cuda.h"
#include <algorithm>
#include <array>
#include <atomic>
#include <charconv>
#include <chrono>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <string>
#include <vector>
#include <unistd.h>

namespace {

const char *kFormatNames[6] = {"P2PKH", "P2PKH (Uncompressed)", "P2WPKH", "P2SH-P2WPKH", "P2TR", "Ethereum"};   // src/address.rs:48-59
const char *kCharsetNames[6] = {"Base58", "Base58", "Bech32", "Base58", "Bech32", "Hex"};                         // src/address.rs:39-45

struct Fail { std::string msg; };
[[noreturn]] void bail(const std::string &m) { throw Fail{m}; }

std::string with_commas(uint64_t n) {                                   // src/lib.rs:1075-1086
    std::string s = std::to_string(n), out;
    for (size_t i = 0; i < s.size(); i++) { if (i && (s.size() - i) % 3 == 0) out.push_back(','); out.push_back(s[i]); }
    return out;
}
std::string fmt_duration(double secs) {                                 // src/lib.rs:1038-1052
    char b[64];
    if (secs < 1.0) snprintf(b, sizeof b, "%.0fms", secs * 1000.0);
    else if (secs < 60.0) snprintf(b, sizeof b, "%.1fs", secs);
    else if (secs < 3600.0) snprintf(b, sizeof b, "%.1fm", secs / 60.0);
    else if (secs < 86400.0) snprintf(b, sizeof b, "%.1fh", secs / 3600.0);
    else if (secs < 31536000.0) snprintf(b, sizeof b, "%.1fd", secs / 86400.0);
    else snprintf(b, sizeof b, "%.1fy", secs / 31536000.0);
    return b;
}
std::string csv_escape(const std::string &f) {                          // src/lib.rs:1058-1073 (RFC 4180)
    if (f.find_first_of(",\"\n\r") == std::string::npos) return f;
    std::string o = "\"";
    for (char c : f) { if (c == '"') o.push_back('"'); o.push_back(c); }
    return o + "\"";
}
std::string json_str(const std::string &s) {
    std::string o = "\"";
    for (unsigned char c : s) {
        if (c == '"' || c == '\\') { o.push_back('\\'); o.push_back((char)c); }
        else if (c == '\n') o += "\\n"; else if (c == '\r') o += "\\r"; else if (c == '\t') o += "\\t";
        else if (c < 0x20) { char b[8]; snprintf(b, sizeof b, "\\u%04x", c); o += b; }
        else o.push_back((char)c);
    }
    return o + "\"";
}
std::string json_f64(double v) {                                        // serde_json: shortest round-trip, always with a fraction or exponent
    char b[64];
    auto r = std::to_chars(b, b + sizeof b, v);
    std::string s(b, r.ptr);
    if (s.find_first_of(".en") == std::string::npos) s += ".0";
    return s;
}

// ---- 256-bit helpers on big-endian byte arrays (range arithmetic of the CLI only) ----
typedef std::array<uint8_t, 32> Key;
Key key_from_hex(const std::string &h, const char *what) {
    if (h.empty() || h.size() > 64) bail(std::string("Invalid ") + what + " hex");
    Key k{}; std::string p(64 - h.size(), '0'); p += h;
    for (int i = 0; i < 32; i++) {
        int v = 0;
        for (int j = 0; j < 2; j++) {
            int c = p[2 * i + j], d;
            if (c >= '0' && c <= '9') d = c - '0'; else if (c >= 'a' && c <= 'f') d = c - 'a' + 10; else if (c >= 'A' && c <= 'F') d = c - 'A' + 10;
            else bail(std::string("Invalid ") + what + " hex");
            v = v * 16 + d;
        }
        k[i] = (uint8_t)v;
    }
    return k;
}
std::string key_hex(const Key &k) { static const char *d = "0123456789abcdef"; std::string s; for (uint8_t b : k) { s.push_back(d[b >> 4]); s.push_back(d[b & 15]); } return s; }
Key key_pow2(unsigned e) { Key k{}; k[31 - e / 8] = (uint8_t)(1u << (e % 8)); return k; }
Key key_add64(const Key &a, uint64_t v, bool *overflow = nullptr) {
    Key r = a; unsigned __int128 c = v;
    for (int i = 31; i >= 0 && c; i--) { c += r[i]; r[i] = (uint8_t)c; c >>= 8; }
    if (overflow) *overflow = c != 0;
    return r;
}
Key key_sub1(const Key &a) { Key r = a; for (int i = 31; i >= 0; i--) { if (r[i]--) break; } return r; }
int key_cmp(const Key &a, const Key &b) { return memcmp(a.data(), b.data(), 32); }
// b - a when it fits 64 bits, else ~0
uint64_t key_diff_sat(const Key &b, const Key &a) {
    Key r{}; int br = 0;
    for (int i = 31; i >= 0; i--) { int d = (int)b[i] - a[i] - br; br = d < 0; r[i] = (uint8_t)(d + (br << 8)); }
    for (int i = 0; i < 24; i++) if (r[i]) return ~0ULL;
    uint64_t v = 0; for (int i = 24; i < 32; i++) v = v << 8 | r[i];
    return v;
}

// ---- argument parsing: --long value | --long=value | -s value | flags ----
struct Args {
    std::map<std::string, std::string> opt;
    bool has(const std::string &k) const { return opt.count(k) != 0; }
    std::string get(const std::string &k, const std::string &def = "") const { auto it = opt.find(k); return it == opt.end() ? def : it->second; }
    uint64_t num(const std::string &k, uint64_t def) const {
        if (!has(k)) return def;
        char *e = nullptr; const std::string v = get(k);
        unsigned long long x = strtoull(v.c_str(), &e, 10);
        if (v.empty() || *e) bail("invalid value '" + v + "' for '--" + k + "'");
        return x;
    }
};
struct OptSpec { const char *lng; char shrt; bool takes_value; };
Args parse_args(int argc, char **argv, int first, const std::vector<OptSpec> &specs) {
    Args a;
    for (int i = first; i < argc; i++) {
        std::string s = argv[i], name, val; bool have_val = false;
        const OptSpec *sp = nullptr;
        if (s.rfind("--", 0) == 0) {
            size_t eq = s.find('=');
            name = s.substr(2, eq == std::string::npos ? std::string::npos : eq - 2);
            if (eq != std::string::npos) { val = s.substr(eq + 1); have_val = true; }
            for (auto &o : specs) if (name == o.lng) sp = &o;
        } else if (s.size() >= 2 && s[0] == '-') {
            for (auto &o : specs) if (o.shrt && s[1] == o.shrt) sp = &o;
            if (sp && s.size() > 2) { val = s.substr(2); have_val = true; }
        }
        if (!sp) bail("unexpected argument '" + s + "'");
        if (sp->takes_value) {
            if (!have_val) { if (i + 1 >= argc) bail(std::string("a value is required for '--") + sp->lng + "'"); val = argv[++i]; }
            a.opt[sp->lng] = val;
        } else a.opt[sp->lng] = "1";
    }
    return a;
}
int parse_format(const std::string &f) {                                 // src/lib.rs:213-242
    static const char *n[6] = {"p2pkh", "p2pkh-uncompressed", "p2wpkh", "p2sh-p2wpkh", "p2tr", "ethereum"};
    for (int i = 0; i < 6; i++) if (f == n[i]) return i;
    bail("invalid value '" + f + "' for '--format': possible values: p2pkh, p2pkh-uncompressed, p2wpkh, p2sh-p2wpkh, p2tr, ethereum");
}

// ---- GPU context ----
struct Runner {
    vgen_ctx *ctx = nullptr;
    std::string device_name;
    ~Runner() { if (ctx) vgen_cuda_destroy(ctx); }
    void open(const Args &a) {
        if (a.has("no-gpu")) bail("--no-gpu: this build has no CPU scan path (the GPU path is the product; see DESIGN.md)");
        std::string be = a.get("backend", "auto");
        std::transform(be.begin(), be.end(), be.begin(), ::tolower);
        if (be != "auto" && be != "cuda") bail("backend '" + be + "' is not available: this build drives B200 GPUs through CUDA only");
        std::vector<int> devs;
        if (a.has("devices")) {
            std::string d = a.get("devices");
            for (size_t p = 0; p <= d.size();) { size_t q = d.find(',', p); if (q == std::string::npos) q = d.size(); devs.push_back(atoi(d.substr(p, q - p).c_str())); p = q + 1; }
        } else devs.push_back(0);
        // the reference's 524,288-key default (src/gpu.rs:83) is tuned for wgpu dispatch; here the default fills the GPU
        uint32_t hint = a.has("gpu-batch-size") ? (uint32_t)a.num("gpu-batch-size", 0) : 0;
        if (vgen_cuda_create(devs.data(), (int)devs.size(), hint, &ctx) != VGEN_OK) bail(std::string("GPU initialization failed: ") + vgen_cuda_last_error(nullptr));
        char name[256] = "";
        vgen_cuda_device_name(devs[0], name, sizeof name);
        device_name = name;
    }
    void check(int rc, const char *what) { if (rc != VGEN_OK) bail(std::string(what) + ": " + vgen_cuda_last_error(ctx)); }
};

std::atomic<int32_t> g_stop{0};
void on_sigint(int) { if (g_stop.exchange(1)) _exit(1); }                 // second Ctrl+C forces exit (src/lib.rs:1088-1098)

struct Match { std::string hex, address, wif; };
struct SearchResult { std::vector<Match> matches; uint64_t operations = 0; double elapsed = 0; };

struct ProgressState { std::chrono::steady_clock::time_point t0, last; uint64_t base = 0; bool json = false, enabled = false; };
void progress_cb(uint64_t keys, void *user) {
    ProgressState *p = (ProgressState *)user;
    if (!p->enabled) return;
    auto now = std::chrono::steady_clock::now();
    if (std::chrono::duration<double>(now - p->last).count() < 1.0) return;
    p->last = now;
    double el = std::chrono::duration<double>(now - p->t0).count();
    uint64_t total = p->base + keys;
    if (p->json) fprintf(stderr, "{\"keys\":%llu,\"elapsed_secs\":%.3f,\"rate\":%.0f}\n", (unsigned long long)total, el, el > 0 ? total / el : 0.0);
    else fprintf(stderr, "\rChecked %s addresses (%.0f Mkeys/s)   ", with_commas(total).c_str(), el > 0 ? total / el / 1e6 : 0.0);
}

// one vgen_cuda_scan over [start, end] (end == nullptr: open-ended), appending formatted matches
void scan_once(Runner &R, const Key &start, const Key *end, uint64_t max_hits, ProgressState &ps, SearchResult &out) {
    uint64_t n_hits = 0, scanned = 0;
    R.check(vgen_cuda_scan(R.ctx, start.data(), end ? end->data() : nullptr, max_hits, (const volatile int32_t *)&g_stop, progress_cb, &ps,
                           nullptr, 0, &n_hits, &scanned), "scan failed");
    out.operations += scanned;
    const uint64_t total = vgen_cuda_scan_total_hits(R.ctx);
    std::vector<char> buf(1 << 20);
    for (uint64_t got = 0; got < total;) {
        uint64_t n = 0;
        int64_t nb = vgen_cuda_scan_fetch_text(R.ctx, got, total - got, buf.data(), buf.size(), &n);
        if (nb <= 0 || n == 0) break;
        const char *p = buf.data(), *e = p + nb;
        while (p < e) {
            const char *nl = (const char *)memchr(p, '\n', e - p); if (!nl) nl = e;
            std::string line(p, nl);
            size_t s1 = line.find(' '), s2 = line.find(' ', s1 + 1);
            if (s1 != std::string::npos && s2 != std::string::npos) out.matches.push_back({line.substr(0, s1), line.substr(s1 + 1, s2 - s1 - 1), line.substr(s2 + 1)});
            p = nl + 1;
        }
        got += n;
    }
}

Key random_valid_key() {                                                  // src/gpu.rs:933-945: random start for `generate`
    Key k{};
    FILE *f = fopen("/dev/urandom", "rb");
    if (!f || fread(k.data(), 1, 32, f) != 32) bail("cannot read /dev/urandom");
    fclose(f);
    k[0] &= 0x7f;                                                         // < 2^255 < n, and nonzero with overwhelming probability
    if (std::all_of(k.begin(), k.end(), [](uint8_t b) { return b == 0; })) k[31] = 1;
    return k;
}

// ---- result writers (src/lib.rs:879-976) ----
void write_results(FILE *w, const std::string &output, const SearchResult &r, const std::string &pattern, int format, bool quiet, bool header) {
    const double rate = r.elapsed > 0 ? r.operations / r.elapsed : 0.0;
    if (output == "csv" && !r.matches.empty() && header) fprintf(w, "address,wif,private_key_hex,format,pattern,operations,elapsed_secs,rate\n");
    for (size_t i = 0; i < r.matches.size(); i++) {
        const Match &m = r.matches[i];
        if (output == "text") {
            fprintf(w, "=== Match %zu of %zu ===\n", i + 1, r.matches.size());
            fprintf(w, "Pattern : %s\nFormat  : %s\nAddress : %s\nWIF     : %s\nHex     : %s\n", pattern.c_str(), kFormatNames[format], m.address.c_str(), m.wif.c_str(), m.hex.c_str());
            if (!quiet) fprintf(w, "Ops     : %s (%.0f/sec)\nTime    : %s\n", with_commas(r.operations).c_str(), rate, fmt_duration(r.elapsed).c_str());
            fprintf(w, "\n");
        } else if (output == "json" || output == "jsonl") {
            const bool pretty = output == "json";
            const char *nl = pretty ? "\n  " : "", *sp = pretty ? " " : "", *end = pretty ? "\n" : "";
            fprintf(w, "{%s\"address\":%s%s,%s\"wif\":%s%s,%s\"private_key_hex\":%s%s,%s\"format\":%s%s,%s\"pattern\":%s%s,%s\"operations\":%s%llu,%s\"elapsed_secs\":%s%s,%s\"rate\":%s%s%s}\n",
                    nl, sp, json_str(m.address).c_str(), nl, sp, json_str(m.wif).c_str(), nl, sp, json_str(m.hex).c_str(), nl, sp, json_str(kFormatNames[format]).c_str(),
                    nl, sp, json_str(pattern).c_str(), nl, sp, (unsigned long long)r.operations, nl, sp, json_f64(r.elapsed).c_str(), nl, sp, json_f64(rate).c_str(), end);
        } else if (output == "csv") {
            fprintf(w, "%s,%s,%s,%s,%s,%llu,%s,%s\n", csv_escape(m.address).c_str(), csv_escape(m.wif).c_str(), csv_escape(m.hex).c_str(), csv_escape(kFormatNames[format]).c_str(),
                    csv_escape(pattern).c_str(), (unsigned long long)r.operations, json_f64(r.elapsed).c_str(), json_f64(rate).c_str());
        } else fprintf(w, "%s\n", m.wif.c_str());                           // minimal
    }
}

// pattern / provider resolution (src/lib.rs:560-640)
struct Resolved { std::string pattern; int format; bool has_range = false; Key start{}, end{}; };
Resolved resolve_pattern(const std::string &pat, const Args &a, int default_format, bool for_range) {
    Resolved r; r.pattern = pat; r.format = default_format;
    char addr[128], err[256] = ""; int fmt = 0, has = 0;
    int rc = vgen_provider_resolve(pat.c_str(), addr, sizeof addr, &fmt, r.start.data(), r.end.data(), &has, err, sizeof err);
    if (rc < 0) bail(err);
    if (rc == 0) {
        if (a.has("prefix-length")) fprintf(stderr, "Warning: --prefix-length is ignored for regex patterns\n");
        return r;
    }
    char out[256];
    if (a.has("prefix-length")) {
        uint64_t len = a.num("prefix-length", 0);
        if (len == 0) bail("--prefix-length must be at least 1 for provider patterns");
        vgen_provider_pattern(addr, (int)std::min<uint64_t>(len, 200), out, sizeof out);
        fprintf(stderr, "Provider: %s \xe2\x86\x92 %s \xe2\x86\x92 pattern '%s'\n", pat.c_str(), addr, out);
    } else {
        vgen_provider_pattern(addr, 0, out, sizeof out);
        if (for_range) fprintf(stderr, "Provider: %s \xe2\x86\x92 %s \xe2\x86\x92 exact match\n", pat.c_str(), addr);
        else fprintf(stderr, "Provider: %s \xe2\x86\x92 %s \xe2\x86\x92 pattern '%s'\n", pat.c_str(), addr, out);
    }
    r.pattern = out; r.format = fmt; r.has_range = has != 0;
    return r;
}

void lint_pattern(const std::string &pattern, bool ci, int format) {      // src/lib.rs:686-708
    char err[256] = "";
    if (vgen_pattern_check(pattern.c_str(), ci, err, sizeof err) != VGEN_OK) bail("Failed to compile pattern: Invalid regex pattern: " + pattern + ": " + err);
    char bad[512] = "";
    if (vgen_pattern_validate_charset(format, pattern.c_str(), ci, bad, sizeof bad) > 0) {
        fprintf(stderr, "Warning: Pattern contains characters not valid in %s addresses: '%s'\n", kCharsetNames[format], bad);
        fprintf(stderr, "  %s alphabet excludes these characters - pattern will NEVER match!\n", kCharsetNames[format]);
        if (format == 0 || format == 1 || format == 3) fprintf(stderr, "  Base58 excludes: 0 (zero), O (uppercase o), I (uppercase i), l (lowercase L)\n");
        fprintf(stderr, "\n");
    }
}

// checkpoint file of a range scan: "vgen-checkpoint 1\nstart=<hex>\nend=<hex>\npattern=<p>\nformat=<n>\nnext=<hex>\nfound=<n>\noperations=<n>\nout_bytes=<n>\n"
// out_bytes: size of the --file output when the checkpoint was written; on resume the file is cut back to it, so that matches
// written after the last checkpoint (a crash between the two writes) are not reported twice when their segment is scanned again
struct Checkpoint { std::string path; Key next{}; uint64_t found = 0, operations = 0, out_bytes = 0; bool resumed = false; };
bool load_checkpoint(Checkpoint &c, const Key &start, const Key &end, const std::string &pattern, int format) {
    std::ifstream f(c.path);
    if (!f) return false;
    std::map<std::string, std::string> kv; std::string line;
    std::getline(f, line);
    if (line != "vgen-checkpoint 1") bail("checkpoint file " + c.path + " is not a vgen checkpoint");
    while (std::getline(f, line)) { size_t eq = line.find('='); if (eq != std::string::npos) kv[line.substr(0, eq)] = line.substr(eq + 1); }
    if (kv["start"] != key_hex(start) || kv["end"] != key_hex(end) || kv["pattern"] != pattern || kv["format"] != std::to_string(format))
        bail("checkpoint file " + c.path + " belongs to a different scan (range, pattern or format differ)");
    c.next = key_from_hex(kv["next"], "checkpoint"); c.found = strtoull(kv["found"].c_str(), nullptr, 10); c.operations = strtoull(kv["operations"].c_str(), nullptr, 10);
    c.out_bytes = strtoull(kv["out_bytes"].c_str(), nullptr, 10);
    c.resumed = true;
    return true;
}
void save_checkpoint(const Checkpoint &c, const Key &start, const Key &end, const std::string &pattern, int format) {
    const std::string tmp = c.path + ".tmp";
    {
        std::ofstream f(tmp, std::ios::trunc);
        f << "vgen-checkpoint 1\nstart=" << key_hex(start) << "\nend=" << key_hex(end) << "\npattern=" << pattern << "\nformat=" << format
          << "\nnext=" << key_hex(c.next) << "\nfound=" << c.found << "\noperations=" << c.operations << "\nout_bytes=" << c.out_bytes << "\n";
        if (!f) bail("cannot write checkpoint file " + tmp);
    }
    if (rename(tmp.c_str(), c.path.c_str()) != 0) bail("cannot replace checkpoint file " + c.path);
}

// unified search runner (src/lib.rs:664-976)
int run_search(const Args &a, const std::string &pattern, bool ci, int format, bool is_range, const Key &start, const Key &end) {
    const bool quiet = a.has("quiet");
    const std::string output = a.get("output", "text");
    if (output != "text" && output != "json" && output != "jsonl" && output != "csv" && output != "minimal")
        bail("invalid value '" + output + "' for '--output': possible values: text, json, jsonl, csv, minimal");
    uint64_t count = a.num("count", 1), repeat = std::max<uint64_t>(1, a.num("repeat", 1));
    if (!is_range && count == 0) bail("--count must be at least 1 for generate");
    lint_pattern(pattern, ci, format);
    Runner R; R.open(a);
    R.check(vgen_cuda_set_pattern(R.ctx, format, pattern.c_str(), ci), "Failed to compile pattern");
    signal(SIGINT, on_sigint);
    FILE *w = stdout;
    if (a.has("file")) { w = fopen(a.get("file").c_str(), a.has("checkpoint") ? "a" : "w"); if (!w) bail("Failed to create output file"); }
    ProgressState ps; ps.json = a.has("progress-json"); ps.enabled = ps.json || (!quiet && isatty(2));
    ps.t0 = ps.last = std::chrono::steady_clock::now();
    if (!quiet) fprintf(stderr, "GPU: %s (CUDA), matcher: %s\n", R.device_name.c_str(), vgen_cuda_matcher_kind(R.ctx));

    SearchResult total;
    auto t0 = std::chrono::steady_clock::now();
    auto elapsed = [&]() { return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); };
    bool wrote_header = false;
    for (uint64_t rep = 0; rep < repeat && !g_stop; rep++) {
        if (!is_range) {
            Key s = random_valid_key();
            ps.base = total.operations;
            scan_once(R, s, nullptr, count, ps, total);
        } else if (!a.has("checkpoint")) {
            ps.base = total.operations;
            scan_once(R, start, &end, count, ps, total);
        } else {
            // resumable scan: segments of --checkpoint-keys keys; after each one its matches are written and the low-water mark saved
            Checkpoint cp; cp.path = a.get("checkpoint"); cp.next = start;
            if (load_checkpoint(cp, start, end, pattern, format)) {
                if (!quiet) fprintf(stderr, "Resuming from %s (%s keys done, %llu match(es) so far)\n", key_hex(cp.next).c_str(), with_commas(cp.operations).c_str(), (unsigned long long)cp.found);
                // drop what was written after the last checkpoint: those keys are at or above cp.next and will be scanned again
                if (w != stdout) {
                    fflush(w); fseek(w, 0, SEEK_END);
                    if ((uint64_t)ftell(w) > cp.out_bytes && ftruncate(fileno(w), (off_t)cp.out_bytes) != 0) bail("cannot truncate the output file to its checkpointed size");
                    fseek(w, 0, SEEK_END);
                }
            }
            const uint64_t seg = std::max<uint64_t>(1ull << 20, a.num("checkpoint-keys", 1ull << 36));
            total.operations = cp.operations;
            bool done = cp.resumed && key_cmp(end, cp.next) < 0;
            while (!done && !g_stop && (count == 0 || cp.found < count)) {
                bool ovf = false;
                Key seg_end = key_add64(cp.next, seg - 1, &ovf);
                if (ovf || key_cmp(end, seg_end) < 0) seg_end = end;
                SearchResult part;
                ps.base = total.operations;
                scan_once(R, cp.next, &seg_end, count ? count - cp.found : 0, ps, part);
                const bool limit_hit = count != 0 && cp.found + part.matches.size() >= count;
                const bool complete = !g_stop && !limit_hit;
                const uint64_t seg_ops = part.operations;
                total.operations += part.operations;
                part.operations = total.operations; part.elapsed = elapsed();
                // order: matches out and flushed first, then the checkpoint (renamed into place) with the new mark and output size
                write_results(w, output, part, pattern, format, quiet, !wrote_header && !cp.resumed);
                wrote_header = wrote_header || !part.matches.empty();
                fflush(w);
                cp.found += part.matches.size();
                total.matches.insert(total.matches.end(), part.matches.begin(), part.matches.end());
                if (complete) { done = key_cmp(seg_end, end) == 0; cp.next = key_add64(seg_end, 1); }
                else if (limit_hit && !part.matches.empty()) cp.next = key_add64(key_from_hex(part.matches.back().hex, "match"), 1);   // hits are the lowest keys: everything up to the last one is done
                else {
                    // interrupted: the library merges windows in key order, so the keys it counted are a prefix of the segment
                    // (scalar 0 is skipped without being counted)
                    const bool from_zero = std::all_of(cp.next.begin(), cp.next.end(), [](uint8_t b) { return b == 0; });
                    cp.next = key_add64(cp.next, seg_ops + (from_zero ? 1 : 0));
                }
                cp.operations = total.operations;
                cp.out_bytes = w != stdout ? (uint64_t)ftell(w) : 0;
                save_checkpoint(cp, start, end, pattern, format);
                if (!complete) break;
            }
        }
    }
    total.elapsed = elapsed();
    if (ps.enabled && !ps.json) fprintf(stderr, "\r%60s\r", "");
    if (!a.has("checkpoint")) write_results(w, output, total, pattern, format, quiet, true);
    if (w != stdout) {
        fclose(w);
        if (!total.matches.empty() && !quiet) fprintf(stderr, "Wrote %zu result(s) to %s\n", total.matches.size(), a.get("file").c_str());
    }
    if (total.matches.empty() && !quiet)
        fprintf(stderr, "No match found after %s operations (%s)\n", with_commas(total.operations).c_str(), fmt_duration(total.elapsed).c_str());
    return 0;
}

void parse_explicit_range(const Args &a, Key &start, Key &end) {          // src/lib.rs:642-663
    if (a.has("puzzle")) {
        uint64_t p = a.num("puzzle", 0);
        if (p < 1 || p > 160) bail("Puzzle number must be between 1 and 160");
        start = key_pow2((unsigned)p - 1); end = key_sub1(key_pow2((unsigned)p));
    } else if (a.has("range")) {
        const std::string r = a.get("range");
        size_t c = r.find(':');
        if (c == std::string::npos || r.find(':', c + 1) != std::string::npos) bail("Range must be in format START:END");
        start = key_from_hex(r.substr(0, c), "start"); end = key_from_hex(r.substr(c + 1), "end");
    } else bail("Either --range, --puzzle, or a provider pattern with key range must be specified");
}

int cmd_estimate(const Args &a) {                                         // src/lib.rs:347-375
    if (!a.has("pattern")) bail("the following required arguments were not provided: --pattern <PATTERN>");
    const bool ci = a.has("ignore-case");
    Resolved r = resolve_pattern(a.get("pattern"), a, parse_format(a.get("format", "p2pkh")), false);
    char err[256] = "";
    if (vgen_pattern_check(r.pattern.c_str(), ci, err, sizeof err) != VGEN_OK) bail("Failed to compile pattern: Invalid regex pattern: " + r.pattern + ": " + err);
    const uint64_t difficulty = vgen_pattern_estimate_difficulty(r.format, r.pattern.c_str(), ci);
    // quick benchmark: the reference times 10,000 CPU addresses; here a few full GPU windows of the scan kernel
    Runner R; R.open(a);
    R.check(vgen_cuda_set_pattern(R.ctx, r.format, r.pattern.c_str(), ci), "Failed to compile pattern");
    Key s = key_pow2(64);
    float ms = 0; uint64_t keys = 0, launches = 0;
    R.check(vgen_cuda_bench_resident(R.ctx, s.data(), 2, &ms, &keys, &launches), "benchmark failed");
    R.check(vgen_cuda_bench_resident(R.ctx, s.data(), 4, &ms, &keys, &launches), "benchmark failed");
    const double rate = keys / (ms * 1e-3);
    printf("Pattern: %s\nFormat: %s\nCase insensitive: %s\n\n", r.pattern.c_str(), kFormatNames[r.format], ci ? "true" : "false");
    printf("Estimated difficulty: 1 in %llu\nBenchmark rate: %.0f addr/sec\nExpected time: %s\n", (unsigned long long)difficulty, rate, fmt_duration(difficulty / rate).c_str());
    return 0;
}

int cmd_verify(const Args &a) {                                           // src/lib.rs:377-494
    if (!a.has("key")) bail("the following required arguments were not provided: --key <KEY>");
    const std::string key = a.get("key");
    Key k{}; int is_wif = 0, compressed = 1;
    int rc = vgen_parse_key(key.c_str(), k.data(), &is_wif, &compressed);
    if (rc == VGEN_ERR_RANGE) bail("malformed or out-of-range secret key");
    if (rc != VGEN_OK) bail("Invalid key format (not WIF or hex)");
    Runner R; R.open(a);
    std::string addr[6];
    for (int f = 0; f < 6; f++) {
        uint8_t dg[32] = {0}; char out[128];
        R.check(vgen_cuda_digest_batch(R.ctx, f, k.data(), 1, dg, f == VGEN_P2TR ? 32 : 20), "address derivation failed");
        if (vgen_format_address(f, dg, out, sizeof out) < 0) bail("address formatting failed");
        addr[f] = out;
    }
    char wif_c[128], wif_u[128];
    vgen_format_wif(VGEN_P2PKH, k.data(), wif_c, sizeof wif_c);
    vgen_format_wif(VGEN_P2PKH_UNCOMPRESSED, k.data(), wif_u, sizeof wif_u);
    printf("Private key: %s\n", is_wif ? key.c_str() : wif_c);
    printf("WIF (uncompr.):     %s\n", wif_u);
    printf("Hex: %s\n\n", key_hex(k).c_str());
    printf("P2PKH address:      %s\n", addr[0].c_str());
    printf("P2PKH (uncompr.):   %s\n", addr[1].c_str());
    printf("P2WPKH address:     %s\n", addr[2].c_str());
    printf("P2SH-P2WPKH addr:  %s\n", addr[3].c_str());
    printf("P2TR address:       %s\n", addr[4].c_str());
    printf("Ethereum address:   %s\n", addr[5].c_str());
    if (a.has("address")) {
        const std::string expected = a.get("address");
        auto lower = [](std::string s) { std::transform(s.begin(), s.end(), s.begin(), ::tolower); return s; };
        const bool is_bech32 = expected.size() >= 3 && lower(expected.substr(0, 3)) == "bc1";
        bool all_lower = true, all_upper = true;
        for (char c : expected) if (isalpha((unsigned char)c)) { all_lower = all_lower && islower((unsigned char)c); all_upper = all_upper && isupper((unsigned char)c); }
        const std::string normalized = (is_bech32 && (all_lower || all_upper)) ? lower(expected) : expected;
        const bool raw_eth = normalized.size() == 40 && std::all_of(normalized.begin(), normalized.end(), [](char c) { return isxdigit((unsigned char)c) != 0; });
        const std::string eth_norm = raw_eth ? "0x" + normalized : normalized;
        if (std::find(addr, addr + 6, normalized) != addr + 6) printf("\nMATCH!\n");
        else if (eth_norm.size() >= 2 && lower(eth_norm.substr(0, 2)) == "0x" && lower(addr[5]) == lower(eth_norm)) printf("\nMATCH! (Ethereum, case-insensitive)\n");
        else printf("\nMISMATCH! Expected: %s\n", expected.c_str());
    }
    return 0;
}

int cmd_list_gpus(const Args &a) {                                        // src/lib.rs:979-1036
    const int n = vgen_cuda_device_count();
    std::vector<std::string> names;
    for (int i = 0; i < n; i++) { char b[256] = ""; vgen_cuda_device_name(i, b, sizeof b); names.push_back(b); }
    if (a.has("json")) {
        printf("[");
        for (int i = 0; i < n; i++)
            printf("%s\n  {\n    \"name\": %s,\n    \"backend\": \"Cuda\",\n    \"device_type\": \"DiscreteGpu\",\n    \"software\": false\n  }", i ? "," : "", json_str(names[i]).c_str());
        printf("%s]\n", n ? "\n" : "");
        return 0;
    }
    printf("Available GPUs:\n");
    if (n == 0) { printf("  (none)\n"); return 0; }
    for (int i = 0; i < n; i++) printf("  %d. %s (Cuda) - DiscreteGpu\n", i + 1, names[i].c_str());
    return 0;
}

const char *kUsage =
    "Bitcoin vanity address generator with regex pattern matching (B200 / CUDA build)\n\n"
    "Usage: vgen <COMMAND>\n\nCommands:\n"
    "  generate   Generate vanity address matching a pattern\n"
    "  estimate   Estimate difficulty of a pattern (dry run)\n"
    "  range      Scan a specific range of keys (e.g., for Bitcoin Puzzles)\n"
    "  verify     Verify a private key produces expected address\n"
    "  list-gpus  List available GPU adapters\n\n"
    "generate: -p PATTERN [-l N] [-f FORMAT] [-i] [-o text|json|jsonl|csv|minimal] [--file PATH] [-c COUNT] [--repeat N] [-q]\n"
    "range:    (-r START:END | --puzzle N | -p boha:b1000:N) [-p PATTERN] [-l N] [-f FORMAT] [-c COUNT (0 = whole range)] [--repeat N]\n"
    "          [-o ...] [--file PATH] [--checkpoint FILE [--checkpoint-keys N]]\n"
    "estimate: -p PATTERN [-l N] [-f FORMAT] [-i]\n"
    "verify:   -k KEY [-a ADDRESS]\n"
    "list-gpus: [--json]\n"
    "common:   [--devices 0,1,...] [--gpu-batch-size N] [--backend auto|cuda] [--progress-json]\n"
    "FORMAT:   p2pkh | p2pkh-uncompressed | p2wpkh | p2sh-p2wpkh | p2tr | ethereum\n";

}  // namespace

int main(int argc, char **argv) {
    if (argc < 2 || !strcmp(argv[1], "-h") || !strcmp(argv[1], "--help") || !strcmp(argv[1], "help")) { fputs(kUsage, argc < 2 ? stderr : stdout); return argc < 2 ? 2 : 0; }
    const std::string cmd = argv[1];
    const std::vector<OptSpec> common = {{"devices", 0, true}, {"gpu-batch-size", 0, true}, {"backend", 0, true}, {"no-gpu", 0, false}, {"progress-json", 0, false},
                                         {"threads", 't', true}, {"cpu-batch-size", 0, true}, {"tui", 0, false}, {"no-tui", 0, false}};
    auto with = [&](std::vector<OptSpec> v) { v.insert(v.end(), common.begin(), common.end()); return v; };
    try {
        if (cmd == "generate") {
            Args a = parse_args(argc, argv, 2, with({{"pattern", 'p', true}, {"prefix-length", 'l', true}, {"format", 'f', true}, {"ignore-case", 'i', false}, {"output", 'o', true},
                                                      {"file", 0, true}, {"count", 'c', true}, {"repeat", 0, true}, {"quiet", 'q', false}}));
            if (!a.has("pattern")) bail("the following required arguments were not provided: --pattern <PATTERN>");
            if (a.has("tui")) fprintf(stderr, "Warning: --tui is deprecated. This build has no TUI; progress goes to stderr.\n");
            const int f = parse_format(a.get("format", "p2pkh"));
            if (a.has("ignore-case") && (f == VGEN_P2WPKH || f == VGEN_ETHEREUM))
                fprintf(stderr, "Warning: Bech32/Ethereum addresses case sensitivity handling is specific. -i flag might be redundant.\n");
            Resolved r = resolve_pattern(a.get("pattern"), a, f, false);
            return run_search(a, r.pattern, a.has("ignore-case"), r.format, false, Key{}, Key{});
        }
        if (cmd == "range") {
            Args a = parse_args(argc, argv, 2, with({{"range", 'r', true}, {"puzzle", 0, true}, {"pattern", 'p', true}, {"prefix-length", 'l', true}, {"format", 'f', true},
                                                      {"output", 'o', true}, {"file", 0, true}, {"count", 'c', true}, {"repeat", 0, true}, {"quiet", 'q', false},
                                                      {"checkpoint", 0, true}, {"checkpoint-keys", 0, true}}));
            Resolved r = resolve_pattern(a.get("pattern", "."), a, parse_format(a.get("format", "p2pkh")), true);
            Key start{}, end{};
            if (a.has("range") || a.has("puzzle")) parse_explicit_range(a, start, end);
            else if (r.has_range) { start = r.start; end = r.end; }
            else if (a.has("pattern") && r.pattern != a.get("pattern")) bail("Provider '" + a.get("pattern") + "' has no key range. Use --range or --puzzle to specify range.");
            else parse_explicit_range(a, start, end);
            if (key_cmp(end, start) < 0) bail("Range end is below range start");      // the reference underflows here (src/scanner.rs:228)
            return run_search(a, r.pattern, false, r.format, true, start, end);
        }
        if (cmd == "estimate") return cmd_estimate(parse_args(argc, argv, 2, with({{"pattern", 'p', true}, {"prefix-length", 'l', true}, {"format", 'f', true}, {"ignore-case", 'i', false}})));
        if (cmd == "verify") return cmd_verify(parse_args(argc, argv, 2, with({{"key", 'k', true}, {"address", 'a', true}})));
        if (cmd == "list-gpus") return cmd_list_gpus(parse_args(argc, argv, 2, with({{"json", 0, false}})));
        fprintf(stderr, "error: unrecognized subcommand '%s'\n\n%s", cmd.c_str(), kUsage);
        return 2;
    } catch (const Fail &f) {
        fprintf(stderr, "Error: %s\n", f.msg.c_str());
        return 1;
    }
}
