// Host-side encoders of the product library (see host_util.h). C++ restatements of the published
// encodings (FIPS 180-4, Keccak, Base58Check, BIP173/BIP350, EIP-55); not shared with oracle/.
#include "host_util.h"
#include <algorithm>
#include <stdlib.h>
#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>
#include <cpuid.h>
#endif

namespace vg {

const char B58_ALPHABET[59] = "123456789ABCDEFGHJKLMNPQRSTUVWXYZabcdefghijkmnopqrstuvwxyz";
const char BECH32_ALPHABET[33] = "qpzry9x8gf2tvdw0s3jn54khce6mua7l";

U256 secp_order() {
    U256 n; n.w[0] = 0xBFD25E8CD0364141ULL; n.w[1] = 0xBAAEDCE6AF48A03BULL; n.w[2] = 0xFFFFFFFFFFFFFFFEULL; n.w[3] = 0xFFFFFFFFFFFFFFFFULL;
    return n;
}
U256 secp_order_minus1() { return secp_order().sub(U256(1)); }

namespace {
inline uint32_t ror(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
const uint32_t K256[64] = {
    0x428a2f98,0x71374491,0xb5c0fbcf,0xe9b5dba5,0x3956c25b,0x59f111f1,0x923f82a4,0xab1c5ed5,0xd807aa98,0x12835b01,0x243185be,0x550c7dc3,
    0x72be5d74,0x80deb1fe,0x9bdc06a7,0xc19bf174,0xe49b69c1,0xefbe4786,0x0fc19dc6,0x240ca1cc,0x2de92c6f,0x4a7484aa,0x5cb0a9dc,0x76f988da,
    0x983e5152,0xa831c66d,0xb00327c8,0xbf597fc7,0xc6e00bf3,0xd5a79147,0x06ca6351,0x14292967,0x27b70a85,0x2e1b2138,0x4d2c6dfc,0x53380d13,
    0x650a7354,0x766a0abb,0x81c2c92e,0x92722c85,0xa2bfe8a1,0xa81a664b,0xc24b8b70,0xc76c51a3,0xd192e819,0xd6990624,0xf40e3585,0x106aa070,
    0x19a4c116,0x1e376c08,0x2748774c,0x34b0bcb5,0x391c0cb3,0x4ed8aa4a,0x5b9cca4f,0x682e6ff3,0x748f82ee,0x78a5636f,0x84c87814,0x8cc70208,
    0x90befffa,0xa4506ceb,0xbef9a3f7,0xc67178f2};
// one compression; rolling 16-word schedule, state rotated by renaming (the confirm workers hash four blocks per hit)
#define VG_S0(x) (ror(x, 2) ^ ror(x, 13) ^ ror(x, 22))
#define VG_S1(x) (ror(x, 6) ^ ror(x, 11) ^ ror(x, 25))
#define VG_s0(x) (ror(x, 7) ^ ror(x, 18) ^ ((x) >> 3))
#define VG_s1(x) (ror(x, 17) ^ ror(x, 19) ^ ((x) >> 10))
#define VG_RND(a, b, c, d, e, f, g, h, i) { \
    uint32_t t1 = h + VG_S1(e) + ((e & f) ^ (~e & g)) + K256[i] + w[(i) & 15]; \
    uint32_t t2 = VG_S0(a) + ((a & b) | (c & (a | b))); \
    d += t1; h = t1 + t2; }
#define VG_SCHED(i) w[(i) & 15] += VG_s0(w[((i) + 1) & 15]) + w[((i) + 9) & 15] + VG_s1(w[((i) + 14) & 15])
void sha_block(uint32_t s[8], const uint8_t *p) {
    uint32_t w[16];
    for (int i = 0; i < 16; i++) w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 | (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
    uint32_t a = s[0], b = s[1], c = s[2], d = s[3], e = s[4], f = s[5], g = s[6], h = s[7];
#pragma GCC unroll 8
    for (int i = 0; i < 64; i += 8) {
        if (i >= 16) { VG_SCHED(i); VG_SCHED(i + 1); VG_SCHED(i + 2); VG_SCHED(i + 3); VG_SCHED(i + 4); VG_SCHED(i + 5); VG_SCHED(i + 6); VG_SCHED(i + 7); }
        VG_RND(a, b, c, d, e, f, g, h, i);     VG_RND(h, a, b, c, d, e, f, g, i + 1); VG_RND(g, h, a, b, c, d, e, f, i + 2); VG_RND(f, g, h, a, b, c, d, e, i + 3);
        VG_RND(e, f, g, h, a, b, c, d, i + 4); VG_RND(d, e, f, g, h, a, b, c, i + 5); VG_RND(c, d, e, f, g, h, a, b, i + 6); VG_RND(b, c, d, e, f, g, h, a, i + 7);
    }
    s[0] += a; s[1] += b; s[2] += c; s[3] += d; s[4] += e; s[5] += f; s[6] += g; s[7] += h;
}
#undef VG_S0
#undef VG_S1
#undef VG_s0
#undef VG_s1
#undef VG_RND
#undef VG_SCHED

// The same compression on the x86 SHA extensions where the host CPU has them (checked once with cpuid; VGEN_NO_SHA_NI=1 keeps the
// portable code): the confirm workers hash four blocks per hit, which is most of their time.
#if defined(__x86_64__) && defined(__GNUC__)
#define VG_HAVE_SHA_NI 1
__attribute__((target("sha,sse4.1,ssse3")))
void sha_block_ni(uint32_t s[8], const uint8_t *p) {
    const __m128i bswap = _mm_set_epi64x(0x0c0d0e0f08090a0bULL, 0x0405060700010203ULL);
    __m128i t = _mm_loadu_si128((const __m128i *)&s[0]);             // D C B A
    __m128i s1 = _mm_loadu_si128((const __m128i *)&s[4]);            // H G F E
    t = _mm_shuffle_epi32(t, 0xB1);                                   // C D A B
    s1 = _mm_shuffle_epi32(s1, 0x1B);                                 // E F G H
    __m128i s0 = _mm_alignr_epi8(t, s1, 8);                           // A B E F
    s1 = _mm_blend_epi16(s1, t, 0xF0);                                // C D G H
    const __m128i save0 = s0, save1 = s1;
    __m128i m[4];
    for (int i = 0; i < 4; i++) m[i] = _mm_shuffle_epi8(_mm_loadu_si128((const __m128i *)(p + 16 * i)), bswap);
    for (int g = 0; g < 16; g++) {                                    // four rounds per step
        __m128i wk = _mm_add_epi32(m[g & 3], _mm_loadu_si128((const __m128i *)&K256[4 * g]));
        s1 = _mm_sha256rnds2_epu32(s1, s0, wk);
        wk = _mm_shuffle_epi32(wk, 0x0E);
        s0 = _mm_sha256rnds2_epu32(s0, s1, wk);
        if (g < 12) {                                                 // W[4g+16 .. 4g+19] takes the place of W[4g .. 4g+3]
            __m128i w = _mm_sha256msg1_epu32(m[g & 3], m[(g + 1) & 3]);
            w = _mm_add_epi32(w, _mm_alignr_epi8(m[(g + 3) & 3], m[(g + 2) & 3], 4));
            m[g & 3] = _mm_sha256msg2_epu32(w, m[(g + 3) & 3]);
        }
    }
    s0 = _mm_add_epi32(s0, save0);
    s1 = _mm_add_epi32(s1, save1);
    t = _mm_shuffle_epi32(s0, 0x1B);                                  // F E B A
    s1 = _mm_shuffle_epi32(s1, 0xB1);                                 // D C H G
    s0 = _mm_blend_epi16(t, s1, 0xF0);                                // D C B A
    s1 = _mm_alignr_epi8(s1, t, 8);                                   // H G F E
    _mm_storeu_si128((__m128i *)&s[0], s0);
    _mm_storeu_si128((__m128i *)&s[4], s1);
}
bool cpu_has_sha_ni() {
    if (getenv("VGEN_NO_SHA_NI")) return false;
    unsigned a = 0, b = 0, c = 0, d = 0;
    if (!__get_cpuid(1, &a, &b, &c, &d) || !(c & (1u << 9)) || !(c & (1u << 19))) return false;      // SSSE3, SSE4.1
    if (!__get_cpuid_count(7, 0, &a, &b, &c, &d)) return false;
    return (b & (1u << 29)) != 0;                                                                     // SHA
}
#endif
inline void sha_compress(uint32_t s[8], const uint8_t *p) {
#ifdef VG_HAVE_SHA_NI
    static const bool ni = cpu_has_sha_ni();
    if (ni) { sha_block_ni(s, p); return; }
#endif
    sha_block(s, p);
}
}  // namespace

void h_sha256(const uint8_t *msg, size_t len, uint8_t out[32]) {
    uint32_t s[8] = {0x6a09e667,0xbb67ae85,0x3c6ef372,0xa54ff53a,0x510e527f,0x9b05688c,0x1f83d9ab,0x5be0cd19};
    size_t o = 0;
    for (; o + 64 <= len; o += 64) sha_compress(s, msg + o);
    uint8_t blk[64];
    size_t r = len - o;
    memcpy(blk, msg + o, r);
    blk[r++] = 0x80;
    if (r > 56) { memset(blk + r, 0, 64 - r); sha_compress(s, blk); r = 0; }
    memset(blk + r, 0, 56 - r);
    const uint64_t bits = (uint64_t)len * 8;
    for (int i = 0; i < 8; i++) blk[56 + i] = (uint8_t)(bits >> (8 * (7 - i)));
    sha_compress(s, blk);
    for (int i = 0; i < 8; i++) { out[4 * i] = s[i] >> 24; out[4 * i + 1] = s[i] >> 16; out[4 * i + 2] = s[i] >> 8; out[4 * i + 3] = s[i]; }
}

namespace {
// Keccak-f[1600] on 25 lanes, lane (x, y) at index x + 5y
void keccak_f(uint64_t st[25]) {
    static const uint64_t RC[24] = {
        0x0000000000000001ULL,0x0000000000008082ULL,0x800000000000808aULL,0x8000000080008000ULL,0x000000000000808bULL,0x0000000080000001ULL,
        0x8000000080008081ULL,0x8000000000008009ULL,0x000000000000008aULL,0x0000000000000088ULL,0x0000000080008009ULL,0x000000008000000aULL,
        0x000000008000808bULL,0x800000000000008bULL,0x8000000000008089ULL,0x8000000000008003ULL,0x8000000000008002ULL,0x8000000000000080ULL,
        0x000000000000800aULL,0x800000008000000aULL,0x8000000080008081ULL,0x8000000000008080ULL,0x0000000080000001ULL,0x8000000080008008ULL};
    // rho rotation of the t-th lane on the pi orbit that starts at (1, 0), and where that lane goes
    static constexpr int ROT[24] = {1, 3, 6, 10, 15, 21, 28, 36, 45, 55, 2, 14, 27, 41, 56, 8, 25, 43, 62, 18, 39, 61, 20, 44};
    static constexpr int DST[24] = {10, 7, 11, 17, 18, 3, 5, 16, 8, 21, 24, 4, 15, 23, 19, 13, 12, 2, 20, 14, 22, 9, 6, 1};
    for (int r = 0; r < 24; r++) {
        uint64_t c[5];
#pragma GCC unroll 5
        for (int x = 0; x < 5; x++) c[x] = st[x] ^ st[x + 5] ^ st[x + 10] ^ st[x + 15] ^ st[x + 20];
#pragma GCC unroll 5
        for (int x = 0; x < 5; x++) {
            const uint64_t n = c[(x + 1) % 5], d = c[(x + 4) % 5] ^ ((n << 1) | (n >> 63));
            st[x] ^= d; st[x + 5] ^= d; st[x + 10] ^= d; st[x + 15] ^= d; st[x + 20] ^= d;
        }
        uint64_t t = st[1];
#pragma GCC unroll 24
        for (int i = 0; i < 24; i++) { const int j = DST[i]; const uint64_t keep = st[j]; st[j] = (t << ROT[i]) | (t >> (64 - ROT[i])); t = keep; }
#pragma GCC unroll 5
        for (int y = 0; y < 25; y += 5) {
            const uint64_t r0 = st[y], r1 = st[y + 1], r2 = st[y + 2], r3 = st[y + 3], r4 = st[y + 4];
            st[y] = r0 ^ (~r1 & r2); st[y + 1] = r1 ^ (~r2 & r3); st[y + 2] = r2 ^ (~r3 & r4); st[y + 3] = r3 ^ (~r4 & r0); st[y + 4] = r4 ^ (~r0 & r1);
        }
        st[0] ^= RC[r];
    }
}
}  // namespace

void h_keccak256(const uint8_t *msg, size_t len, uint8_t out[32]) {
    const size_t rate = 136;
    uint64_t st[25] = {};
    auto absorb = [&](const uint8_t *blk) {
        for (size_t i = 0; i < rate / 8; i++) { uint64_t v = 0; for (int b = 7; b >= 0; b--) v = (v << 8) | blk[8 * i + b]; st[i] ^= v; }
        keccak_f(st);
    };
    size_t o = 0;
    for (; o + rate <= len; o += rate) absorb(msg + o);
    uint8_t blk[136];
    memset(blk, 0, sizeof blk);
    memcpy(blk, msg + o, len - o);
    blk[len - o] ^= 0x01; blk[rate - 1] ^= 0x80;       // original Keccak padding (not SHA-3's 0x06)
    absorb(blk);
    for (int i = 0; i < 4; i++) for (int b = 0; b < 8; b++) out[8 * i + b] = (uint8_t)(st[i] >> (8 * b));
}

// Base58 of a big-endian byte string. The number is divided by 58^5 on 32-bit limbs (one 64-by-32 division per limb and
// five digits per pass) instead of by 58 on bytes: ~100 divisions for a 25-byte address payload instead of ~850.
std::string h_base58(const uint8_t *data, size_t len) {
    size_t zeros = 0;
    while (zeros < len && data[zeros] == 0) zeros++;
    const size_t nb = len - zeros;
    std::vector<uint32_t> limb((nb + 3) / 4, 0);      // most significant limb first
    for (size_t i = 0; i < nb; i++) {
        const size_t from_end = nb - 1 - i;            // byte significance
        limb[limb.size() - 1 - from_end / 4] |= (uint32_t)data[zeros + i] << (8 * (from_end % 4));
    }
    std::string digits;                                // least significant first
    digits.reserve(nb * 138 / 100 + 8);
    size_t first = 0;
    while (first < limb.size() && limb[first] == 0) first++;
    while (first < limb.size()) {
        uint64_t rem = 0;
        for (size_t i = first; i < limb.size(); i++) { const uint64_t cur = (rem << 32) | limb[i]; limb[i] = (uint32_t)(cur / 656356768u); rem = cur % 656356768u; }
        while (first < limb.size() && limb[first] == 0) first++;
        const bool last = first == limb.size();        // the most significant group carries no leading zeros
        for (int k = 0; k < 5 && (!last || rem); k++) { digits.push_back(B58_ALPHABET[rem % 58]); rem /= 58; }
    }
    std::string out(zeros, '1');
    out.append(digits.rbegin(), digits.rend());
    return out;
}
std::string h_base58check(const uint8_t *payload, size_t len) {
    uint8_t h1[32], h2[32];
    h_sha256(payload, len, h1); h_sha256(h1, 32, h2);
    if (len <= 60) {
        uint8_t buf[64];
        memcpy(buf, payload, len); memcpy(buf + len, h2, 4);
        return h_base58(buf, len + 4);
    }
    std::vector<uint8_t> buf(payload, payload + len);
    buf.insert(buf.end(), h2, h2 + 4);
    return h_base58(buf.data(), buf.size());
}
bool h_base58check_decode(const std::string &s, std::vector<uint8_t> &payload) {
    size_t zeros = 0;
    while (zeros < s.size() && s[zeros] == '1') zeros++;
    std::vector<uint8_t> num;   // little-endian base-256
    for (size_t i = zeros; i < s.size(); i++) {
        const char *q = strchr(B58_ALPHABET, s[i]);
        if (!q || !s[i]) return false;
        unsigned carry = (unsigned)(q - B58_ALPHABET);
        for (auto &b : num) { unsigned cur = b * 58u + carry; b = (uint8_t)cur; carry = cur >> 8; }
        while (carry) { num.push_back((uint8_t)carry); carry >>= 8; }
        if (num.size() > 64) return false;
    }
    std::vector<uint8_t> full(zeros, 0);
    full.insert(full.end(), num.rbegin(), num.rend());
    if (full.size() < 5) return false;
    uint8_t h1[32], h2[32];
    h_sha256(full.data(), full.size() - 4, h1); h_sha256(h1, 32, h2);
    if (memcmp(h2, full.data() + full.size() - 4, 4) != 0) return false;
    payload.assign(full.begin(), full.end() - 4);
    return true;
}
namespace {
uint32_t polymod_step(uint32_t c) {
    static const uint32_t G[5] = {0x3b6a57b2u, 0x26508e6du, 0x1ea119fau, 0x3d4233ddu, 0x2a1462b3u};
    uint32_t top = c >> 25; c = (c & 0x1ffffffu) << 5;
    for (int i = 0; i < 5; i++) c ^= (0u - ((top >> i) & 1u)) & G[i];
    return c;
}
uint32_t bech32_checksum_state(const std::vector<uint8_t> &data5) {
    uint32_t c = 1;
    const char hrp[] = "bc";
    for (int i = 0; i < 2; i++) c = polymod_step(c) ^ (uint32_t)(hrp[i] >> 5);
    c = polymod_step(c);
    for (int i = 0; i < 2; i++) c = polymod_step(c) ^ (uint32_t)(hrp[i] & 31);
    for (uint8_t d : data5) c = polymod_step(c) ^ d;
    return c;
}
}  // namespace
std::string h_segwit(int witver, const uint8_t *prog, size_t len) {
    uint8_t d5[80]; size_t n = 0;                     // witness programs are at most 40 bytes: 1 + 64 groups + 6 checksum symbols
    if (len > 40) len = 40;
    d5[n++] = (uint8_t)witver;
    uint32_t acc = 0; int bits = 0;
    for (size_t i = 0; i < len; i++) { acc = (acc << 8) | prog[i]; bits += 8; while (bits >= 5) { bits -= 5; d5[n++] = (acc >> bits) & 31; } }
    if (bits) d5[n++] = (acc << (5 - bits)) & 31;
    uint32_t c = 1;
    const char hrp[] = "bc";
    for (int i = 0; i < 2; i++) c = polymod_step(c) ^ (uint32_t)(hrp[i] >> 5);
    c = polymod_step(c);
    for (int i = 0; i < 2; i++) c = polymod_step(c) ^ (uint32_t)(hrp[i] & 31);
    for (size_t i = 0; i < n; i++) c = polymod_step(c) ^ d5[i];
    for (int i = 0; i < 6; i++) c = polymod_step(c);
    const uint32_t pm = c ^ (witver == 0 ? 1u : 0x2bc830a3u);
    std::string out;
    out.reserve(3 + n + 6);
    out = "bc1";
    for (size_t i = 0; i < n; i++) out.push_back(BECH32_ALPHABET[d5[i]]);
    for (int i = 0; i < 6; i++) out.push_back(BECH32_ALPHABET[(pm >> (5 * (5 - i))) & 31]);
    return out;
}
bool h_segwit_decode(const std::string &addr, int &witver, std::vector<uint8_t> &prog) {
    if (addr.size() < 11 || addr.compare(0, 3, "bc1") != 0) return false;
    std::vector<uint8_t> d5;
    for (size_t i = 3; i < addr.size(); i++) { const char *q = strchr(BECH32_ALPHABET, addr[i]); if (!q || !addr[i]) return false; d5.push_back((uint8_t)(q - BECH32_ALPHABET)); }
    uint32_t pm = bech32_checksum_state(d5);
    witver = d5[0];
    if (pm != (witver == 0 ? 1u : 0x2bc830a3u)) return false;
    prog.clear();
    uint32_t acc = 0; int bits = 0;
    for (size_t i = 1; i + 6 < d5.size(); i++) { acc = (acc << 5) | d5[i]; bits += 5; if (bits >= 8) { bits -= 8; prog.push_back((uint8_t)(acc >> bits)); } }
    if (bits >= 5 || (acc & ((1u << bits) - 1))) return false;
    return true;
}
std::string h_hex(const uint8_t *d, size_t len) {
    static const char H[] = "0123456789abcdef";
    std::string s; s.reserve(len * 2);
    for (size_t i = 0; i < len; i++) { s.push_back(H[d[i] >> 4]); s.push_back(H[d[i] & 15]); }
    return s;
}
std::string h_eip55(const uint8_t addr[20]) {
    std::string low = h_hex(addr, 20);
    uint8_t h[32]; h_keccak256((const uint8_t *)low.data(), 40, h);
    std::string out = "0x";
    for (int i = 0; i < 40; i++) {
        int nib = (i & 1) ? (h[i / 2] & 15) : (h[i / 2] >> 4);
        char c = low[i];
        out.push_back((c >= 'a' && nib >= 8) ? (char)(c - 32) : c);
    }
    return out;
}
int digest_len_of_format(int format) { return format == 4 ? 32 : 20; }

std::string format_address(int format, const uint8_t *digest) {
    uint8_t payload[21];
    switch (format) {
        case 0: case 1: payload[0] = 0x00; memcpy(payload + 1, digest, 20); return h_base58check(payload, 21);
        case 3: payload[0] = 0x05; memcpy(payload + 1, digest, 20); return h_base58check(payload, 21);
        case 2: return h_segwit(0, digest, 20);
        case 4: return h_segwit(1, digest, 32);
        case 5: return h_eip55(digest);
    }
    return std::string();
}
std::string format_wif(int format, const uint8_t key_be[32]) {
    if (format == 5) return h_hex(key_be, 32);
    uint8_t p[34]; p[0] = 0x80; memcpy(p + 1, key_be, 32); p[33] = 0x01;
    return h_base58check(p, format == 1 ? 33 : 34);
}
}  // namespace vg
