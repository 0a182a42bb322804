// Pattern lowering: (format, regex) -> device matcher.
//   ^literal...   -> inclusive intervals over the leading 64 digest bits (no per-key Base58/Bech32 encode)
//   ^address$     -> exact digest compare (the address is decoded once on the host)
//   anything else -> dense DFA over the on-GPU address encode
// Every device hit is re-checked on the host with the full DFA over the re-encoded address string, so
// interval bounds only need to be supersets (they are exact up to the 4 Base58Check checksum bytes).
// Reference semantics: src/pattern.rs:21-45 (match = regex search on the address string),
// src/provider.rs:55-63 (prefix / exact pattern shapes), src/address.rs:92-151 (formats).
#include "lowering.h"
#include <algorithm>
#include <string.h>

namespace vg {
namespace {

const char *alphabet_for(int format) {
    switch (format) {
        case VG_FMT_P2PKH: case VG_FMT_P2PKH_UNCOMPRESSED: case VG_FMT_P2SH_P2WPKH: return B58_ALPHABET;
        case VG_FMT_P2WPKH: case VG_FMT_P2TR: return "qpzry9x8gf2tvdw0s3jn54khce6mua7lb1";   // data charset + hrp/separator chars
        default: return "0123456789abcdefABCDEFx";
    }
}
typedef std::pair<U256, U256> Range;   // inclusive range over the digest read as a big-endian integer of `bits` bits

U256 pow58(unsigned e) { U256 r(1); for (unsigned i = 0; i < e; i++) r = r.mul64(58); return r; }
U256 umax(const U256 &a, const U256 &b) { return a < b ? b : a; }
U256 umin(const U256 &a, const U256 &b) { return a < b ? a : b; }

// Base58Check payload = version || h160 || checksum4 (25 bytes)
void base58_ranges(int version, const std::string &p, std::vector<Range> &out) {
    size_t a = 0;
    while (a < p.size() && p[a] == '1') a++;
    std::string rest = p.substr(a);
    const U256 max160 = U256::pow2(160).sub(U256(1));
    if (rest.empty()) {
        if (a == 0) { out.push_back({U256(0), max160}); return; }
        if (version != 0) return;
        // at least a leading zero bytes in the 25-byte payload: h160||ck < 2^(8*(25-a))
        if (a > 25) return;
        unsigned bits = 8 * (25 - (unsigned)a);
        U256 hi = bits >= 32 ? U256::pow2(bits).sub(U256(1)).shr(32) : U256(0);
        out.push_back({U256(0), umin(hi, max160)});
        return;
    }
    U256 zlo, zhi, verbase;
    if (version == 0) {
        if (a < 1 || a > 24) return;
        zlo = U256::pow2(8 * (24 - (unsigned)a)); zhi = U256::pow2(8 * (25 - (unsigned)a)).sub(U256(1));
    } else {
        if (a != 0) return;
        verbase = U256((uint64_t)version).shl(192);
        zlo = verbase; zhi = U256((uint64_t)version + 1).shl(192).sub(U256(1));
    }
    U256 val; size_t r = rest.size();
    for (char c : rest) {
        const char *q = strchr(B58_ALPHABET, c);
        if (!q || !c) return;
        val = val.mul64(58).add64((uint64_t)(q - B58_ALPHABET));
    }
    for (unsigned D = (unsigned)r; D <= 35; D++) {
        U256 dlo = pow58(D - 1), dhi = pow58(D).sub(U256(1));
        U256 scale = pow58(D - (unsigned)r);
        uint64_t ov1 = 0;
        U256 plo = val;
        // val * 58^(D-r): multiply stepwise to detect overflow
        bool overflow = false;
        for (unsigned i = 0; i < D - r; i++) { plo = plo.mul64(58, &ov1); if (ov1) { overflow = true; break; } }
        if (overflow) break;
        bool carry = false;
        U256 phi = plo.add(scale.sub(U256(1)), &carry);
        if (carry) phi = U256(0).sub(U256(1));   // saturate at 2^256 - 1
        U256 lo = umax(umax(dlo, zlo), plo), hi = umin(umin(dhi, zhi), phi);
        if (hi < lo) continue;
        out.push_back({lo.sub(verbase).shr(32), hi.sub(verbase).shr(32)});
    }
}
// bech32 data characters after "bc1<q|p>" map 5 bits each onto the top of the witness program
void bech32_ranges(char verchar, unsigned prog_bits, const std::string &p, std::vector<Range> &out) {
    const std::string head = std::string("bc1") + verchar;
    const U256 maxv = U256::pow2(prog_bits).sub(U256(1));
    size_t n = std::min(p.size(), head.size());
    if (p.compare(0, n, head, 0, n) != 0) return;
    if (p.size() <= head.size()) { out.push_back({U256(0), maxv}); return; }
    unsigned max_chars = prog_bits / 5;   // whole characters that lie inside the program
    U256 bits; unsigned r = 0;
    for (size_t i = head.size(); i < p.size() && r < max_chars; i++, r++) {
        const char *q = strchr(BECH32_ALPHABET, p[i]);
        if (!q || !p[i]) return;
        bits = bits.shl(5).add64((uint64_t)(q - BECH32_ALPHABET));
    }
    unsigned sh = prog_bits - 5 * r;
    out.push_back({bits.shl(sh), bits.add64(1).shl(sh).sub(U256(1))});
}
void hex_ranges(const std::string &p, std::vector<Range> &out) {
    const std::string head = "0x";
    const U256 maxv = U256::pow2(160).sub(U256(1));
    size_t n = std::min(p.size(), head.size());
    if (p.compare(0, n, head, 0, n) != 0) return;
    if (p.size() <= 2) { out.push_back({U256(0), maxv}); return; }
    U256 bits; unsigned r = 0;
    for (size_t i = 2; i < p.size() && r < 40; i++, r++) {
        int c = p[i], v;
        if (c >= '0' && c <= '9') v = c - '0';
        else if (c >= 'a' && c <= 'f') v = c - 'a' + 10;
        else if (c >= 'A' && c <= 'F') v = c - 'A' + 10;
        else return;
        bits = bits.shl(4).add64((uint64_t)v);
    }
    unsigned sh = 160 - 4 * r;
    out.push_back({bits.shl(sh), bits.add64(1).shl(sh).sub(U256(1))});
}
void ranges_for(int format, const std::string &p, std::vector<Range> &out) {
    switch (format) {
        case VG_FMT_P2PKH: case VG_FMT_P2PKH_UNCOMPRESSED: base58_ranges(0, p, out); break;
        case VG_FMT_P2SH_P2WPKH: base58_ranges(5, p, out); break;
        case VG_FMT_P2WPKH: bech32_ranges('q', 160, p, out); break;
        case VG_FMT_P2TR: bech32_ranges('p', 256, p, out); break;
        default: hex_ranges(p, out); break;
    }
}
// exact address -> digest; false if the string is not a valid address of this format
bool decode_exact(int format, const std::string &addr, uint8_t digest[32]) {
    std::vector<uint8_t> payload; int witver;
    memset(digest, 0, 32);
    switch (format) {
        case VG_FMT_P2PKH: case VG_FMT_P2PKH_UNCOMPRESSED: case VG_FMT_P2SH_P2WPKH:
            if (!h_base58check_decode(addr, payload) || payload.size() != 21) return false;
            if (payload[0] != (format == VG_FMT_P2SH_P2WPKH ? 5 : 0)) return false;
            if (h_base58check(payload.data(), 21) != addr) return false;   // canonical form only
            memcpy(digest, payload.data() + 1, 20); return true;
        case VG_FMT_P2WPKH:
            if (!h_segwit_decode(addr, witver, payload) || witver != 0 || payload.size() != 20) return false;
            memcpy(digest, payload.data(), 20); return true;
        case VG_FMT_P2TR:
            if (!h_segwit_decode(addr, witver, payload) || witver != 1 || payload.size() != 32) return false;
            memcpy(digest, payload.data(), 32); return true;
        default: {
            if (addr.size() != 42 || addr[0] != '0' || addr[1] != 'x') return false;
            for (int i = 0; i < 20; i++) {
                int v = 0;
                for (int j = 0; j < 2; j++) {
                    int c = addr[2 + 2 * i + j], d;
                    if (c >= '0' && c <= '9') d = c - '0'; else if (c >= 'a' && c <= 'f') d = c - 'a' + 10; else if (c >= 'A' && c <= 'F') d = c - 'A' + 10; else return false;
                    v = v * 16 + d;
                }
                digest[i] = (uint8_t)v;
            }
            return h_eip55(digest) == addr;   // the reference matches the EIP-55 form, case-sensitively
        }
    }
}
// device view of the DFA for `format`: alphabet-index -> class map, start state advanced over the fixed head of the
// address, and whether hex-letter case matters (Ethereum)
void fill_device_dfa(int format, const Dfa &dfa, vg_dfa_dev &d) {
    memset(&d, 0, sizeof d);
    d.n_states = dfa.n_states; d.n_classes = dfa.n_classes; d.start = dfa.start;
    d.table_u16 = dfa.n_states * dfa.n_classes; d.accept_bytes = dfa.n_states;
    auto adv = [&](const char *head) { for (const char *c = head; *c; c++) d.start = dfa.table[(size_t)d.start * dfa.n_classes + dfa.class_of[(uint8_t)*c]]; };
    switch (format) {
        case VG_FMT_P2PKH: case VG_FMT_P2PKH_UNCOMPRESSED: case VG_FMT_P2SH_P2WPKH:
            for (int i = 0; i < 58; i++) d.sym_class[i] = dfa.class_of[(uint8_t)B58_ALPHABET[i]];
            break;
        case VG_FMT_P2WPKH: case VG_FMT_P2TR:
            for (int i = 0; i < 32; i++) d.sym_class[i] = dfa.class_of[(uint8_t)BECH32_ALPHABET[i]];
            adv(format == VG_FMT_P2WPKH ? "bc1q" : "bc1p");
            break;
        default:
            for (int i = 0; i < 16; i++) {
                d.sym_class[i] = dfa.class_of[(uint8_t)"0123456789abcdef"[i]];
                d.sym_class[16 + i] = dfa.class_of[(uint8_t)"0123456789ABCDEF"[i]];
                if (d.sym_class[i] != d.sym_class[16 + i]) d.flags |= 1u;
            }
            adv("0x");
            break;
    }
}

// Literal prefixes read off the DFA itself, for anchored patterns whose syntax tree is not a plain run of character sets
// (alternations such as ^(1Cat|1Dog), optional groups, ...): breadth-first over the format's alphabet from the start
// state, dropping transitions into dead states (no accepting state reachable) and closing a prefix when its state is
// universal (everything reachable accepts). Stops growing when more than 64 prefixes would be live; what is returned is
// then a necessary condition only (`complete` = false) and the host confirmation decides.
bool derive_prefixes_from_dfa(const Dfa &d, const char *alpha, std::vector<std::string> &prefixes, bool &complete) {
    const uint32_t ns = d.n_states, nc = d.n_classes;
    std::vector<uint8_t> live(ns, 0), universal(ns, 1);
    for (uint32_t s = 0; s < ns; s++) live[s] = d.accept_end[s] != 0;
    for (bool ch = true; ch;) {                 // live: an accepting state is reachable
        ch = false;
        for (uint32_t s = 0; s < ns; s++) if (!live[s]) for (uint32_t c = 0; c < nc; c++) if (live[d.table[(size_t)s * nc + c]]) { live[s] = 1; ch = true; break; }
    }
    for (uint32_t s = 0; s < ns; s++) universal[s] = d.accept_end[s] != 0;
    for (bool ch = true; ch;) {                 // universal: accepting, and every successor universal
        ch = false;
        for (uint32_t s = 0; s < ns; s++) if (universal[s]) for (uint32_t c = 0; c < nc; c++) if (!universal[d.table[(size_t)s * nc + c]]) { universal[s] = 0; ch = true; break; }
    }
    if (!live[d.start]) { prefixes.clear(); complete = true; return true; }     // can never match
    std::string chars;
    for (const char *a = alpha; *a; a++) if (chars.find(*a) == std::string::npos) chars.push_back(*a);
    std::vector<std::pair<uint32_t, std::string>> frontier{{d.start, ""}};
    std::vector<std::string> closed;
    complete = true;
    for (int depth = 0; depth < 64 && !frontier.empty(); depth++) {
        std::vector<std::pair<uint32_t, std::string>> next;
        std::vector<std::string> closed_next = closed;
        for (auto &f : frontier) {
            if (universal[f.first]) { closed_next.push_back(f.second); continue; }
            if (d.accept_end[f.first]) { closed_next.push_back(f.second); complete = false; }     // the address may END here ("...$")
            for (char c : chars) {
                uint32_t t = d.table[(size_t)f.first * nc + d.class_of[(uint8_t)c]];
                if (live[t]) next.push_back({t, f.second + c});
            }
        }
        if (closed_next.size() + next.size() > 64) { complete = false; break; }
        closed.swap(closed_next);
        frontier.swap(next);
    }
    prefixes = closed;
    for (auto &f : frontier) {
        prefixes.push_back(f.second);
        if (!universal[f.first]) complete = false;      // e.g. "...$": the string must END here; the prefix alone does not decide
    }
    return !(prefixes.size() == 1 && prefixes[0].empty());
}
}  // namespace

uint32_t dfa_smem_bytes(const vg_dfa_dev &d) {
    uint64_t tw = ((uint64_t)d.table_u16 + 1) / 2, aw = ((uint64_t)d.accept_bytes + 3) / 4;
    uint64_t bytes = (tw + aw + 16) * 4;
    return bytes <= 64 * 1024 ? (uint32_t)bytes : 0;
}

// prefixes -> merged inclusive intervals over the leading 64 digest bits -> device matcher (kind stays DFA when useless)
// `decides`: the prefixes are the whole pattern (nothing follows them), so an unconstrained prefix means "match all"
static bool intervals_from_prefixes(int format, const std::vector<std::string> &prefixes, bool decides, LoweredPattern &out) {
    std::vector<Range> ranges;
    for (const auto &p : prefixes) ranges_for(format, p, ranges);
    // reduce to the leading 64 digest bits, merge
    unsigned bits = format == VG_FMT_P2TR ? 256 : 160;
    std::vector<std::pair<uint64_t, uint64_t>> r64;
    for (auto &r : ranges) r64.push_back({r.first.shr(bits - 64).w[0], r.second.shr(bits - 64).w[0]});
    std::sort(r64.begin(), r64.end());
    std::vector<std::pair<uint64_t, uint64_t>> merged;
    for (auto &r : r64) {
        if (!merged.empty() && (r.first <= merged.back().second || r.first == merged.back().second + 1)) merged.back().second = std::max(merged.back().second, r.second);
        else merged.push_back(r);
    }
    if (merged.empty()) { out.dev.kind = VG_MATCH_INTERVALS; out.dev.n_intervals = 0; out.kind_name = "none"; out.density = 0.0; return true; }   // e.g. ^3 on P2PKH
    if (merged.size() > VG_MAX_INTERVALS) return true;   // too fragmented: keep the DFA path
    if (merged.size() == 1 && merged[0].first == 0 && merged[0].second == ~0ULL) {
        // the prefix constrains nothing (e.g. ^1 on P2PKH, ^bc1q on P2WPKH): only worth a prefilter if the rest does
        if (decides) { out.dev.kind = VG_MATCH_ALL; out.kind_name = "all"; }
        return true;
    }
    out.dev.kind = VG_MATCH_INTERVALS; out.kind_name = "intervals";
    out.dev.n_intervals = (int)merged.size();
    out.density = 0.0;
    for (auto &m : merged) out.density += ((double)(m.second - m.first) + 1.0) / 18446744073709551616.0;
    for (size_t i = 0; i < merged.size(); i++) {
        out.dev.lo_hi[i][0] = (uint32_t)(merged[i].first >> 32); out.dev.lo_hi[i][1] = (uint32_t)merged[i].first;
        out.dev.lo_hi[i][2] = (uint32_t)(merged[i].second >> 32); out.dev.lo_hi[i][3] = (uint32_t)merged[i].second;
    }
    // The intervals are only a prefilter when something follows the prefix (^1C.*z$, ^bc1qa.*zz$) or when the pattern tells
    // a-f from A-F (Ethereum ^0xAb: the intervals ignore the EIP-55 case): the device then runs the staged DFA on the
    // survivors, so that only real matches cross PCIe. Without it every prefix candidate (1.3 M records per 77.6 M-key window
    // for a 2-character Base58 prefix) would be re-encoded on the host and the scan would be host-bound.
    if (!out.host_regex && (!decides || (format == VG_FMT_ETHEREUM && (out.dev.dfa.flags & 1u)))) out.dev.confirm_dfa = 1;
    return true;
}

static bool lower_syntactic(int format, const std::string &pattern, bool ci, LoweredPattern &out, std::string &err) {
    out = LoweredPattern();
    out.format = format; out.pattern = pattern; out.case_insensitive = ci;
    if (format < 0 || format > 5) { err = "invalid address format"; return false; }
    if (!compile_pattern(pattern, ci, out.dfa, out.info, err)) { err = "Invalid regex pattern: " + pattern + ": " + err; return false; }   // src/pattern.rs:32-33
    memset(&out.dev, 0, sizeof out.dev);
    fill_device_dfa(format, out.dfa, out.dev.dfa);
    // A pattern whose DFA is too large for a device table (regex_dfa.cpp): the device reports a superset - every key, or the keys
    // inside the intervals of a literal prefix when the pattern has one - and the host decides with the NFA. Exact, host-bound.
    out.host_regex = !out.dfa.table_complete;
    if (format == VG_FMT_ETHEREUM && (out.dev.dfa.flags & 1u) && !out.host_regex) {
        // the pattern tells a-f from A-F, so the device needs the EIP-55 checksum hash of a key before it can run the DFA.
        // Append the same pattern compiled case-insensitively: it accepts the lower-case hex of every address the pattern
        // matches, needs no hash, and rejects almost every key (measured: F00$ 1,785 -> see profiles/configs_r01_c.log).
        Dfa fold; PatternInfo fi; std::string ferr;
        if (compile_pattern(pattern, true, fold, fi, ferr, true) && out.dfa.table.size() + fold.table.size() < 60000) {
            vg_dfa_dev fd;
            fill_device_dfa(format, fold, fd);
            vg_dfa_dev &d = out.dev.dfa;
            d.fold_table = (uint32_t)out.dfa.table.size(); d.fold_accept = (uint32_t)out.dfa.accept_end.size();
            d.fold_classes = fold.n_classes; d.fold_start = fd.start;
            memcpy(d.fold_sym, fd.sym_class, 16);
            out.dfa.table.insert(out.dfa.table.end(), fold.table.begin(), fold.table.end());
            out.dfa.accept_end.insert(out.dfa.accept_end.end(), fold.accept_end.begin(), fold.accept_end.end());
            d.table_u16 = (uint32_t)out.dfa.table.size(); d.accept_bytes = (uint32_t)out.dfa.accept_end.size();
        }
    }
    if (out.host_regex) {
        out.dev.kind = VG_MATCH_ALL; out.kind_name = "all";
        const PatternInfo &hi = out.info;
        if (!hi.anchored_start || hi.prefix_sets.empty()) return true;
    } else
    if (format == VG_FMT_P2WPKH || format == VG_FMT_P2TR) {
        // "dead$": a literal of 1..6 Bech32 characters anchored at the end lies inside the checksum symbols
        static const char *charset = "qpzry9x8gf2tvdw0s3jn54khce6mua7l";
        const size_t n = pattern.size();
        if (n >= 2 && n <= 7 && pattern[n - 1] == '$') {
            uint32_t value = 0; bool ok = true;
            for (size_t i = 0; i + 1 < n && ok; i++) {
                char c = pattern[i];
                if (ci && c >= 'A' && c <= 'Z') c = (char)(c - 'A' + 'a');
                const char *q = strchr(charset, c);
                if (c == 0 || !q) ok = false; else value = (value << 5) | (uint32_t)(q - charset);
            }
            if (ok) { out.dev.dfa.chk_mask = (1u << (5 * (n - 1))) - 1u; out.dev.dfa.chk_value = value; }
        }
    }
    if (!out.host_regex && (format == VG_FMT_P2PKH || format == VG_FMT_P2PKH_UNCOMPRESSED || format == VG_FMT_P2SH_P2WPKH)) {
        // "Cat$": a literal of 1..5 Base58 characters anchored at the end = the payload modulo 58^k (up to 8 residues when
        // the pattern is case-insensitive: every spelling the alphabet has)
        const size_t n = pattern.size();
        if (n >= 2 && n <= 6 && pattern[n - 1] == '$') {
            std::vector<uint32_t> vals{0};
            bool ok = true;
            for (size_t i = 0; i + 1 < n && ok; i++) {
                const char c = pattern[i];
                std::string spell(1, c);
                if (ci && c >= 'a' && c <= 'z') spell.push_back((char)(c - 32));
                if (ci && c >= 'A' && c <= 'Z') spell.push_back((char)(c + 32));
                std::vector<uint32_t> nx;
                const bool alnum = (c >= '0' && c <= '9') || (c >= 'a' && c <= 'z') || (c >= 'A' && c <= 'Z');
                for (char sc : spell) {
                    const char *q = alnum ? strchr(B58_ALPHABET, sc) : nullptr;
                    if (!q) continue;
                    for (uint32_t v : vals) nx.push_back(v * 58u + (uint32_t)(q - B58_ALPHABET));
                }
                if (!alnum || nx.size() > 8) ok = false;        // a metacharacter, or too many spellings: no prefilter
                vals.swap(nx);                                   // empty: no spelling is Base58 -> the DFA rejects every key anyway
            }
            if (ok && !vals.empty()) {
                vg_dfa_dev &d = out.dev.dfa;
                uint64_t M = 1;
                for (size_t i = 0; i + 1 < n; i++) M *= 58;
                d.b58_mod = (uint32_t)M; d.b58_nvals = (uint32_t)vals.size();
                for (size_t i = 0; i < vals.size(); i++) d.b58_vals[i] = vals[i];
                uint64_t w = 1 % M;
                for (int j = 0; j < 13; j++) { d.b58_w[j] = (uint32_t)w; w = (w << 16) % M; }
                const uint64_t mu = (uint64_t)((((unsigned __int128)1) << 64) / M);
                d.b58_mu_lo = (uint32_t)mu; d.b58_mu_hi = (uint32_t)(mu >> 32);
            }
        }
    }
    if (!out.host_regex && format == VG_FMT_ETHEREUM) {
        // "F00$", "dead$": a literal of 1..8 hex digits anchored at the end. Whatever the EIP-55 case turns out to be, the
        // lower-case hex of a matching address ends with the lower-cased literal: a test on the last digest bits.
        const size_t n = pattern.size();
        if (n >= 2 && n <= 9 && pattern[n - 1] == '$') {
            uint32_t value = 0; bool ok = true;
            for (size_t i = 0; i + 1 < n && ok; i++) {
                const char c = pattern[i];
                uint32_t d = (c >= '0' && c <= '9') ? (uint32_t)(c - '0') : (c >= 'a' && c <= 'f') ? (uint32_t)(c - 'a' + 10)
                           : (c >= 'A' && c <= 'F') ? (uint32_t)(c - 'A' + 10) : 16u;
                if (d > 15u) ok = false; else value = (value << 4) | d;
            }
            if (ok) { out.dev.dfa.chk_mask = n - 1 == 8 ? 0xffffffffu : (1u << (4 * (n - 1))) - 1u; out.dev.dfa.chk_value = value; }
        }
    }
    if (!out.host_regex) { out.dev.kind = VG_MATCH_DFA; out.kind_name = "dfa"; }
    const PatternInfo &info = out.info;
    if (!info.anchored_start || info.prefix_sets.empty()) return true;

    // expand the leading sets into literal prefixes (restricted to characters this format can produce)
    const char *alpha = alphabet_for(format);
    std::vector<std::string> prefixes{""};
    bool truncated = false;
    for (const auto &set : info.prefix_sets) {
        std::string chars;
        for (const char *a = alpha; *a; a++) if (set[(unsigned char)*a] && chars.find(*a) == std::string::npos) chars.push_back(*a);
        if (chars.empty()) { out.dev.kind = VG_MATCH_INTERVALS; out.dev.n_intervals = 0; out.kind_name = "none"; out.density = 0.0; return true; }   // can never match
        if (prefixes.size() * chars.size() > 64) { truncated = true; break; }
        std::vector<std::string> nx;
        for (const auto &p : prefixes) for (char c : chars) nx.push_back(p + c);
        prefixes.swap(nx);
    }
    if (prefixes.size() == 1 && prefixes[0].empty()) return true;

    if (!truncated && info.rest_is_end && prefixes.size() == 1) {
        uint8_t dg[32];
        if (!decode_exact(format, prefixes[0], dg)) { out.dev.kind = VG_MATCH_INTERVALS; out.dev.n_intervals = 0; out.kind_name = "none"; out.density = 0.0; return true; }
        out.dev.kind = VG_MATCH_EXACT; out.kind_name = "exact"; out.density = 0.0;
        out.dev.exact_words = digest_len_of_format(format) / 4;
        for (int i = 0; i < 8; i++) out.dev.exact[i] = (uint32_t)dg[4 * i] | (uint32_t)dg[4 * i + 1] << 8 | (uint32_t)dg[4 * i + 2] << 16 | (uint32_t)dg[4 * i + 3] << 24;
        return true;
    }
    return intervals_from_prefixes(format, prefixes, info.rest_empty && !truncated, out);
}

bool lower_pattern(int format, const std::string &pattern, bool ci, LoweredPattern &out, std::string &err) {
    if (!lower_syntactic(format, pattern, ci, out, err)) return false;
    if (out.dev.kind != VG_MATCH_DFA || !out.info.anchored_start || out.host_regex) return true;
    // anchored, but the leading run of character sets decided nothing (alternations such as ^(1Cat|1Dog), ^bc1q(aa|zz),
    // optional groups): take the prefixes from the DFA. Superset semantics: the host confirmation decides.
    std::vector<std::string> pf; bool complete = false;
    if (!derive_prefixes_from_dfa(out.dfa, alphabet_for(format), pf, complete)) return true;
    if (pf.empty()) { out.dev.kind = VG_MATCH_INTERVALS; out.dev.n_intervals = 0; out.kind_name = "none"; out.density = 0.0; return true; }
    return intervals_from_prefixes(format, pf, complete, out);
}

}  // namespace vg
