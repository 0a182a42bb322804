// Pattern lint, difficulty heuristic and the provider shim used by the CLI.
// Behaviour follows the reference: Pattern::validate_charset (src/pattern.rs:48-175), Pattern::estimate_difficulty
// (src/pattern.rs:177-253), count_fixed_chars (src/pattern.rs:266-293), provider::resolve / build_pattern /
// build_exact_pattern (src/provider.rs:12-63). Pure host code; nothing here touches the GPU.
#include "pattern_lint.h"
#include <algorithm>
#include <string.h>
#include <ctype.h>

namespace vg {
namespace {
const char *valid_chars_for(int format) {
    switch (format) {
        case 0: case 1: case 3: return "123456789ABCDEFGHJKLMNPQRSTUVWXYZabcdefghijkmnopqrstuvwxyz";
        case 2: case 4: return "023456789acdefghjklmnpqrstuvwxyz";
        default: return "0123456789abcdefABCDEFx";
    }
}
bool contains(const std::string &s, char c) { return s.find(c) != std::string::npos; }
void push_unique(std::string &s, char c) { if (!contains(s, c)) s.push_back(c); }
bool is_alnum(char c) { return isalnum((unsigned char)c) != 0; }   // patterns are ASCII on this path
}  // namespace

std::string validate_charset(const std::string &pattern, bool ci, int format) {
    const std::string valid = valid_chars_for(format);
    auto is_valid = [&](char c) {
        if (ci) return contains(valid, (char)tolower((unsigned char)c)) || contains(valid, (char)toupper((unsigned char)c));
        return contains(valid, c);
    };
    std::string invalid, cls;
    bool in_class = false, negated = false, class_start = false, escaped = false, pending_range = false, have_prev = false;
    char prev = 0;
    for (char c : pattern) {
        if (escaped) {
            escaped = false;
            if (in_class) { class_start = false; push_unique(cls, c); }
            continue;
        }
        if (c == '\\') { escaped = true; continue; }
        if (c == '[') { in_class = true; class_start = true; negated = false; cls.clear(); have_prev = false; pending_range = false; continue; }
        if (c == ']' && in_class) {
            if (!negated) {
                bool has_valid = false;
                for (char ch : cls) has_valid = has_valid || is_valid(ch);
                if (!has_valid) for (char ch : cls) push_unique(invalid, ch);
            }
            in_class = false; have_prev = false; pending_range = false;
            continue;
        }
        if (c == '^' && in_class && class_start) { negated = true; class_start = false; continue; }
        if (!in_class && strchr("^$.*+?(){}|", c)) { class_start = false; continue; }
        if (c == '-' && in_class) { class_start = false; if (have_prev) pending_range = true; continue; }
        if (is_alnum(c)) {
            class_start = false;
            if (in_class) {
                if (pending_range) {
                    if (have_prev) {
                        unsigned char from = (unsigned char)std::min(prev, c), to = (unsigned char)std::max(prev, c);
                        for (unsigned ch = from; ch <= to; ch++) push_unique(cls, (char)ch);
                    } else push_unique(cls, c);
                    pending_range = false;
                } else push_unique(cls, c);
                prev = c; have_prev = true;
            } else if (!is_valid(c)) push_unique(invalid, c);
            continue;
        }
        class_start = false;
        if (in_class) push_unique(cls, c);
    }
    return invalid;
}

size_t count_fixed_chars(const std::string &pattern) {
    size_t count = 0;
    bool in_class = false, escaped = false;
    for (char c : pattern) {
        if (escaped) { escaped = false; continue; }
        if (c == '\\') escaped = true;
        else if (c == '[') in_class = true;
        else if (c == ']') in_class = false;
        else if (strchr("^$.*+?(){}|", c)) {}
        else if (!in_class && is_alnum(c)) count++;
    }
    return count;
}

static uint64_t sat_pow(uint64_t b, unsigned e) {
    uint64_t r = 1;
    for (unsigned i = 0; i < e; i++) {
        if (r > ~0ULL / b) return ~0ULL;
        r *= b;
    }
    return r;
}

uint64_t estimate_difficulty(const std::string &pattern, bool ci, int format) {
    uint64_t alphabet = (format == 0 || format == 1 || format == 3) ? (ci ? 34 : 58) : (format == 5 ? 16 : 32);
    size_t fixed = count_fixed_chars(pattern), sub = 0;
    if (!pattern.empty() && pattern[0] == '^') {
        const std::string p = pattern.substr(1);
        auto sw = [&](const char *s) { return p.compare(0, strlen(s), s) == 0; };
        switch (format) {
            case 0: case 1: sub = sw("1"); break;
            case 3: sub = sw("3"); break;
            case 2: sub = sw("bc1q") ? 4 : sw("bc1") ? 3 : sw("bc") ? 2 : sw("b"); break;
            case 4: sub = sw("bc1p") ? 4 : sw("bc1") ? 3 : sw("bc") ? 2 : sw("b"); break;
            default: sub = (sw("0x") || sw("0X")) ? 2 : sw("0"); break;
        }
    }
    size_t eff = fixed > sub ? fixed - sub : 0;
    if (eff == 0) return 1;
    return sat_pow(alphabet, (unsigned)eff);
}

// ---- provider shim (src/provider.rs:12-63) ------------------------------------------------------------------------
// The reference resolves "boha:<collection>:<id>" through the boha crate (Cargo.lock: boha 0.11.0), whose puzzle data
// is not available offline. Embedded here: the b1000 entries whose address is public knowledge and pinned by tests
// (#1..#20: solved keys re-derived by the oracle in tests/test_host.py; #66: README.md:107-108). Anything else is an
// error that names the gap instead of guessing.
namespace {
struct PuzzleEntry { int id; const char *address; };
const PuzzleEntry kB1000[] = {
    {1, "1BgGZ9tcN4rm9KBzDn7KprQz87SZ26SAMH"}, {2, "1CUNEBjYrCn2y1SdiUMohaKUi4wpP326Lb"}, {3, "19ZewH8Kk1PDbSNdJ97FP4EiCjTRaZMZQA"},
    {4, "1EhqbyUMvvs7BfL8goY6qcPbD6YKfPqb7e"}, {5, "1E6NuFjCi27W5zoXg8TRdcSRq84zJeBW3k"}, {6, "1PitScNLyp2HCygzadCh7FveTnfmpPbfp8"},
    {7, "1McVt1vMtCC7yn5b9wgX1833yCcLXzueeC"}, {8, "1M92tSqNmQLYw33fuBvjmeadirh1ysMBxK"}, {9, "1CQFwcjw1dwhtkVWBttNLDtqL7ivBonGPV"},
    {10, "1LeBZP5QCwwgXRtmVUvTVrraqPUokyLHqe"}, {11, "1PgQVLmst3Z314JrQn5TNiys8Hc38TcXJu"}, {12, "1DBaumZxUkM4qMQRt2LVWyFJq5kDtSZQot"},
    {13, "1Pie8JkxBT6MGPz9Nvi3fsPkr2D8q3GBc1"}, {14, "1ErZWg5cFCe4Vw5BzgfzB74VNLaXEiEkhk"}, {15, "1QCbW9HWnwQWiQqVo5exhAnmfqKRrCRsvW"},
    {16, "1BDyrQ6WoF8VN3g9SAS1iKZcPzFfnDVieY"}, {17, "1HduPEXZRdG26SUT5Yk83mLkPyjnZuJ7Bm"}, {18, "1GnNTmTVLZiqQfLbAdp9DVdicEnB5GoERE"},
    {19, "1NWmZRpHH4XSPwsW6dsS3nrNWfL1yrJj4w"}, {20, "1HsMJxNiV7TLxmoF6uJNkydxPFDog4NQum"}, {66, "13zb1hQbWVsc2S7ZTZnP2G4undNNpdh5so"},
};
}  // namespace

int provider_resolve(const std::string &pattern, ProviderResult &out, std::string &err) {
    size_t colon = pattern.find(':');
    if (colon == std::string::npos) return 0;
    if (pattern.substr(0, colon) != "boha") return 0;                       // unknown providers are plain regexes
    std::string path = pattern.substr(colon + 1);
    std::replace(path.begin(), path.end(), ':', '/');
    if (path.compare(0, 6, "b1000/") != 0) { err = "Failed to get puzzle '" + path + "': only the b1000 collection is embedded in this build"; return -1; }
    const std::string id = path.substr(6);
    if (id.empty() || id.size() > 3 || !std::all_of(id.begin(), id.end(), [](char c) { return c >= '0' && c <= '9'; })) {
        err = "Failed to get puzzle '" + path + "': invalid puzzle number"; return -1;
    }
    const int n = atoi(id.c_str());
    for (const PuzzleEntry &e : kB1000) {
        if (e.id != n) continue;
        out.address = e.address; out.format = 0; out.has_range = true;
        out.start = U256::pow2((unsigned)n - 1); out.end = U256::pow2((unsigned)n).sub(U256(1));      // 2^(n-1) .. 2^n - 1
        return 1;
    }
    err = "Failed to get puzzle '" + path + "': not in the embedded table (the boha data set is not available offline)";
    return -1;
}

static std::string regex_escape(const std::string &s) {                    // regex::escape
    std::string o;
    for (char c : s) { if (strchr("\\.+*?()|[]{}^$#&-~", c)) o.push_back('\\'); o.push_back(c); }
    return o;
}
std::string provider_build_pattern(const ProviderResult &r, size_t prefix_len) {
    return "^" + regex_escape(r.address.substr(0, std::min(prefix_len, r.address.size())));
}
std::string provider_build_exact_pattern(const ProviderResult &r) { return "^" + regex_escape(r.address) + "$"; }

}  // namespace vg
