// Host-side encoders of the product library (see host_util.h). C++ restatements of the published
// encodings (FIPS 180-4, Keccak, Base58Check, BIP173/BIP350, EIP-55); not shared with oracle/.
#include "host_util.h"
#include <algorithm>

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
void sha_block(uint32_t s[8], const uint8_t *p) {
    uint32_t w[64];
    for (int i = 0; i < 16; i++) w[i] = (uint32_t)p[4 * i] << 24 | (uint32_t)p[4 * i + 1] << 16 | (uint32_t)p[4 * i + 2] << 8 | p[4 * i + 3];
    for (int i = 16; i < 64; i++)
        w[i] = w[i - 16] + (ror(w[i - 15], 7) ^ ror(w[i - 15], 18) ^ (w[i - 15] >> 3)) + w[i - 7] + (ror(w[i - 2], 17) ^ ror(w[i - 2], 19) ^ (w[i - 2] >> 10));
    uint32_t v[8]; memcpy(v, s, 32);
    for (int i = 0; i < 64; i++) {
        uint32_t t1 = v[7] + (ror(v[4], 6) ^ ror(v[4], 11) ^ ror(v[4], 25)) + ((v[4] & v[5]) ^ (~v[4] & v[6])) + K256[i] + w[i];
        uint32_t t2 = (ror(v[0], 2) ^ ror(v[0], 13) ^ ror(v[0], 22)) + ((v[0] & v[1]) ^ (v[0] & v[2]) ^ (v[1] & v[2]));
        for (int j = 7; j > 0; j--) v[j] = v[j - 1];
        v[4] += t1; v[0] = t1 + t2;
    }
    for (int i = 0; i < 8; i++) s[i] += v[i];
}
}  // namespace

void h_sha256(const uint8_t *msg, size_t len, uint8_t out[32]) {
    uint32_t s[8] = {0x6a09e667,0xbb67ae85,0x3c6ef372,0xa54ff53a,0x510e527f,0x9b05688c,0x1f83d9ab,0x5be0cd19};
    std::vector<uint8_t> m(msg, msg + len);
    m.push_back(0x80);
    while (m.size() % 64 != 56) m.push_back(0);
    uint64_t bits = (uint64_t)len * 8;
    for (int i = 7; i >= 0; i--) m.push_back((uint8_t)(bits >> (8 * i)));
    for (size_t o = 0; o < m.size(); o += 64) sha_block(s, m.data() + o);
    for (int i = 0; i < 8; i++) { out[4 * i] = s[i] >> 24; out[4 * i + 1] = s[i] >> 16; out[4 * i + 2] = s[i] >> 8; out[4 * i + 3] = s[i]; }
}

void h_keccak256(const uint8_t *msg, size_t len, uint8_t out[32]) {
    static const uint64_t RC[24] = {
        0x0000000000000001ULL,0x0000000000008082ULL,0x800000000000808aULL,0x8000000080008000ULL,0x000000000000808bULL,0x0000000080000001ULL,
        0x8000000080008081ULL,0x8000000000008009ULL,0x000000000000008aULL,0x0000000000000088ULL,0x0000000080008009ULL,0x000000008000000aULL,
        0x000000008000808bULL,0x800000000000008bULL,0x8000000000008089ULL,0x8000000000008003ULL,0x8000000000008002ULL,0x8000000000000080ULL,
        0x000000000000800aULL,0x800000008000000aULL,0x8000000080008081ULL,0x8000000000008080ULL,0x0000000080000001ULL,0x8000000080008008ULL};
    uint64_t a[5][5] = {};   // a[x][y]
    const size_t rate = 136;
    std::vector<uint8_t> m(msg, msg + len);
    size_t pad = rate - m.size() % rate;
    m.resize(m.size() + pad, 0);
    m[len] ^= 0x01; m[m.size() - 1] ^= 0x80;
    auto rol = [](uint64_t v, unsigned n) { return n ? (v << n) | (v >> (64 - n)) : v; };
    for (size_t o = 0; o < m.size(); o += rate) {
        for (size_t i = 0; i < rate / 8; i++) { uint64_t v = 0; for (int b = 7; b >= 0; b--) v = (v << 8) | m[o + 8 * i + b]; a[i % 5][i / 5] ^= v; }
        for (int r = 0; r < 24; r++) {
            uint64_t c[5], d[5], b[5][5];
            for (int x = 0; x < 5; x++) c[x] = a[x][0] ^ a[x][1] ^ a[x][2] ^ a[x][3] ^ a[x][4];
            for (int x = 0; x < 5; x++) d[x] = c[(x + 4) % 5] ^ rol(c[(x + 1) % 5], 1);
            for (int x = 0; x < 5; x++) for (int y = 0; y < 5; y++) a[x][y] ^= d[x];
            // rho + pi: walk the (x, y) -> (y, 2x+3y) orbit with triangular-number rotations
            b[0][0] = a[0][0];
            int x = 1, y = 0;
            for (int t = 0; t < 24; t++) {
                unsigned rot = ((t + 1) * (t + 2) / 2) % 64;
                b[y][(2 * x + 3 * y) % 5] = rol(a[x][y], rot);
                int nx = y, ny = (2 * x + 3 * y) % 5; x = nx; y = ny;
            }
            for (int xx = 0; xx < 5; xx++) for (int yy = 0; yy < 5; yy++) a[xx][yy] = b[xx][yy] ^ (~b[(xx + 1) % 5][yy] & b[(xx + 2) % 5][yy]);
            a[0][0] ^= RC[r];
        }
    }
    for (int i = 0; i < 4; i++) for (int b = 0; b < 8; b++) out[8 * i + b] = (uint8_t)(a[i % 5][i / 5] >> (8 * b));
}

std::string h_base58(const uint8_t *data, size_t len) {
    size_t zeros = 0;
    while (zeros < len && data[zeros] == 0) zeros++;
    std::vector<uint8_t> num(data + zeros, data + len);   // big-endian base-256
    std::string digits;
    size_t start = 0;
    while (start < num.size()) {
        unsigned rem = 0;
        for (size_t i = start; i < num.size(); i++) { unsigned cur = rem * 256 + num[i]; num[i] = (uint8_t)(cur / 58); rem = cur % 58; }
        digits.push_back(B58_ALPHABET[rem]);
        while (start < num.size() && num[start] == 0) start++;
    }
    std::string out(zeros, '1');
    out.append(digits.rbegin(), digits.rend());
    return out;
}
std::string h_base58check(const uint8_t *payload, size_t len) {
    uint8_t h1[32], h2[32];
    h_sha256(payload, len, h1); h_sha256(h1, 32, h2);
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
    for (int i = 0; i < 5; i++) if ((top >> i) & 1) c ^= G[i];
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
    std::vector<uint8_t> d5; d5.push_back((uint8_t)witver);
    uint32_t acc = 0; int bits = 0;
    for (size_t i = 0; i < len; i++) { acc = (acc << 8) | prog[i]; bits += 8; while (bits >= 5) { bits -= 5; d5.push_back((acc >> bits) & 31); } }
    if (bits) d5.push_back((acc << (5 - bits)) & 31);
    std::vector<uint8_t> ext = d5; ext.insert(ext.end(), 6, 0);
    uint32_t pm = bech32_checksum_state(ext) ^ (witver == 0 ? 1u : 0x2bc830a3u);
    std::string out = "bc1";
    for (uint8_t d : d5) out.push_back(BECH32_ALPHABET[d]);
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
