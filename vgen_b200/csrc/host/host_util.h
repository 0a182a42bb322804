// Host-side helpers of the product library: 256-bit integers, and the string encoders the reference
// runs on the host for every key (src/gpu.rs:1034-1060 address strings, :1080 WIF; src/address.rs:
// 176-198 EIP-55). Here they run only for hits. Independent of oracle/ (which is test-only).
#pragma once
#include <stdint.h>
#include <string.h>
#include <string>
#include <vector>
#include <array>

namespace vg {

struct U256 {
    uint64_t w[4];   // little-endian
    U256() : w{0, 0, 0, 0} {}
    explicit U256(uint64_t v) : w{v, 0, 0, 0} {}
    static U256 from_be(const uint8_t b[32]) {
        U256 r;
        for (int i = 0; i < 4; i++) { uint64_t v = 0; for (int j = 0; j < 8; j++) v = (v << 8) | b[(3 - i) * 8 + j]; r.w[i] = v; }
        return r;
    }
    void to_be(uint8_t b[32]) const {
        for (int i = 0; i < 4; i++) for (int j = 0; j < 8; j++) b[(3 - i) * 8 + j] = (uint8_t)(w[i] >> (56 - 8 * j));
    }
    void to_limbs32(uint32_t l[8]) const { for (int i = 0; i < 4; i++) { l[2 * i] = (uint32_t)w[i]; l[2 * i + 1] = (uint32_t)(w[i] >> 32); } }
    bool is_zero() const { return (w[0] | w[1] | w[2] | w[3]) == 0; }
    int cmp(const U256 &o) const { for (int i = 3; i >= 0; i--) { if (w[i] < o.w[i]) return -1; if (w[i] > o.w[i]) return 1; } return 0; }
    bool operator<(const U256 &o) const { return cmp(o) < 0; }
    bool operator<=(const U256 &o) const { return cmp(o) <= 0; }
    bool operator==(const U256 &o) const { return cmp(o) == 0; }
    U256 add(const U256 &o, bool *carry = nullptr) const {
        U256 r; unsigned __int128 c = 0;
        for (int i = 0; i < 4; i++) { c += (unsigned __int128)w[i] + o.w[i]; r.w[i] = (uint64_t)c; c >>= 64; }
        if (carry) *carry = c != 0;
        return r;
    }
    U256 add64(uint64_t v, bool *carry = nullptr) const { return add(U256(v), carry); }
    U256 sub(const U256 &o, bool *borrow = nullptr) const {
        U256 r; uint64_t br = 0;
        for (int i = 0; i < 4; i++) { unsigned __int128 t = (unsigned __int128)w[i] - o.w[i] - br; r.w[i] = (uint64_t)t; br = (uint64_t)(t >> 64) & 1; }
        if (borrow) *borrow = br != 0;
        return r;
    }
    U256 mul64(uint64_t v, uint64_t *overflow = nullptr) const {
        U256 r; unsigned __int128 c = 0;
        for (int i = 0; i < 4; i++) { c += (unsigned __int128)w[i] * v; r.w[i] = (uint64_t)c; c >>= 64; }
        if (overflow) *overflow = (uint64_t)c;
        return r;
    }
    uint64_t divmod64(uint64_t d) {   // in-place divide, returns the remainder
        unsigned __int128 rem = 0;
        for (int i = 3; i >= 0; i--) { unsigned __int128 cur = (rem << 64) | w[i]; w[i] = (uint64_t)(cur / d); rem = cur % d; }
        return (uint64_t)rem;
    }
    U256 shr(unsigned n) const {
        U256 r; unsigned q = n / 64, s = n % 64;
        for (int i = 0; i < 4; i++) {
            uint64_t lo = (i + q < 4) ? w[i + q] : 0, hi = (i + q + 1 < 4) ? w[i + q + 1] : 0;
            r.w[i] = s ? (lo >> s) | (hi << (64 - s)) : lo;
        }
        return r;
    }
    U256 shl(unsigned n) const {
        U256 r; unsigned q = n / 64, s = n % 64;
        for (int i = 3; i >= 0; i--) {
            uint64_t hi = (i >= (int)q) ? w[i - q] : 0, lo = (i >= (int)q + 1) ? w[i - q - 1] : 0;
            r.w[i] = s ? (hi << s) | (lo >> (64 - s)) : hi;
        }
        return r;
    }
    bool fits64() const { return (w[1] | w[2] | w[3]) == 0; }
    static U256 pow2(unsigned n) { return U256(1).shl(n); }
};
U256 secp_order();          // n
U256 secp_order_minus1();   // n - 1

void h_sha256(const uint8_t *msg, size_t len, uint8_t out[32]);
void h_keccak256(const uint8_t *msg, size_t len, uint8_t out[32]);
std::string h_base58(const uint8_t *data, size_t len);
std::string h_base58check(const uint8_t *payload, size_t len);
bool h_base58check_decode(const std::string &s, std::vector<uint8_t> &payload);   // verifies the checksum
std::string h_segwit(int witver, const uint8_t *prog, size_t len);
bool h_segwit_decode(const std::string &addr, int &witver, std::vector<uint8_t> &prog);   // "bc" only, checks the checksum
std::string h_eip55(const uint8_t addr[20]);
std::string h_hex(const uint8_t *d, size_t len);
int digest_len_of_format(int format);

// address string of a digest (format enum as in include/vgen_cuda.h); empty string for a bad format
std::string format_address(int format, const uint8_t *digest);
// WIF of a secret (hex for Ethereum, src/address.rs:110; uncompressed flag src/address.rs:116-118,146)
std::string format_wif(int format, const uint8_t key_be[32]);

extern const char B58_ALPHABET[59];
extern const char BECH32_ALPHABET[33];
}  // namespace vg
