/* ORACLE (test infrastructure, not product): Base58Check (base58ck 0.1.0), Bech32/Bech32m
 * (bech32 0.11.1; BIP173/BIP350), EIP-55 (src/address.rs:176-198), WIF (src/address.rs:146).
 * Plain byte-at-a-time restatements of the published encodings. */
#ifndef VGEN_ORACLE_ENCODE_H
#define VGEN_ORACLE_ENCODE_H
#include "hashes.h"

static const char B58_ALPHABET[] = "123456789ABCDEFGHJKLMNPQRSTUVWXYZabcdefghijkmnopqrstuvwxyz";
static const char BECH32_CHARSET[] = "qpzry9x8gf2tvdw0s3jn54khce6mua7l";

/* out must hold len*138/100+2 bytes; returns string length */
static inline int base58_encode(const uint8_t *in, int len, char *out) {
    uint8_t digits[128]; int nd = 0, zeros = 0;
    while (zeros < len && in[zeros] == 0) zeros++;
    for (int i = zeros; i < len; i++) {
        int carry = in[i];
        for (int j = 0; j < nd; j++) { carry += digits[j] << 8; digits[j] = carry % 58; carry /= 58; }
        while (carry) { digits[nd++] = carry % 58; carry /= 58; }
    }
    int o = 0;
    for (int i = 0; i < zeros; i++) out[o++] = '1';
    for (int i = nd - 1; i >= 0; i--) out[o++] = B58_ALPHABET[digits[i]];
    out[o] = 0; return o;
}
static inline int base58check_encode(const uint8_t *payload, int len, char *out) {
    uint8_t buf[96], ck[32];
    memcpy(buf, payload, len); sha256d(payload, len, ck); memcpy(buf + len, ck, 4);
    return base58_encode(buf, len + 4, out);
}
/* returns decoded length or -1 */
static inline int base58_decode(const char *s, uint8_t *out, int cap) {
    uint8_t bytes[128]; int nb = 0, zeros = 0; size_t n = strlen(s);
    while ((size_t)zeros < n && s[zeros] == '1') zeros++;
    for (size_t i = zeros; i < n; i++) {
        const char *q = strchr(B58_ALPHABET, s[i]); if (!q || !s[i]) return -1;
        int carry = (int)(q - B58_ALPHABET);
        for (int j = 0; j < nb; j++) { carry += bytes[j] * 58; bytes[j] = carry & 0xff; carry >>= 8; }
        while (carry) { if (nb >= 120) return -1; bytes[nb++] = carry & 0xff; carry >>= 8; }
    }
    if (zeros + nb > cap) return -1;
    int o = 0;
    for (int i = 0; i < zeros; i++) out[o++] = 0;
    for (int i = nb - 1; i >= 0; i--) out[o++] = bytes[i];
    return o;
}
static inline uint32_t bech32_polymod_step(uint32_t pre) {
    uint32_t b = pre >> 25;
    return ((pre & 0x1FFFFFF) << 5) ^ (-((b >> 0) & 1) & 0x3b6a57b2UL) ^ (-((b >> 1) & 1) & 0x26508e6dUL) ^
           (-((b >> 2) & 1) & 0x1ea119faUL) ^ (-((b >> 3) & 1) & 0x3d4233ddUL) ^ (-((b >> 4) & 1) & 0x2a1462b3UL);
}
/* segwit address: hrp "bc", witness version 0 (bech32) or 1 (bech32m), program of plen bytes */
static inline int segwit_encode(const char *hrp, int witver, const uint8_t *prog, int plen, char *out) {
    uint8_t data[80]; int nd = 0;
    data[nd++] = (uint8_t)witver;
    uint32_t acc = 0; int bits = 0;
    for (int i = 0; i < plen; i++) {
        acc = (acc << 8) | prog[i]; bits += 8;
        while (bits >= 5) { bits -= 5; data[nd++] = (acc >> bits) & 31; }
    }
    if (bits) data[nd++] = (acc << (5 - bits)) & 31;
    uint32_t chk = 1; size_t hl = strlen(hrp);
    for (size_t i = 0; i < hl; i++) chk = bech32_polymod_step(chk) ^ (hrp[i] >> 5);
    chk = bech32_polymod_step(chk);
    for (size_t i = 0; i < hl; i++) chk = bech32_polymod_step(chk) ^ (hrp[i] & 31);
    for (int i = 0; i < nd; i++) chk = bech32_polymod_step(chk) ^ data[i];
    for (int i = 0; i < 6; i++) chk = bech32_polymod_step(chk);
    chk ^= (witver == 0) ? 1 : 0x2bc830a3UL;
    int o = 0;
    for (size_t i = 0; i < hl; i++) out[o++] = hrp[i];
    out[o++] = '1';
    for (int i = 0; i < nd; i++) out[o++] = BECH32_CHARSET[data[i]];
    for (int i = 0; i < 6; i++) out[o++] = BECH32_CHARSET[(chk >> (5 * (5 - i))) & 31];
    out[o] = 0; return o;
}
static inline void hex_encode(const uint8_t *in, int len, char *out) {
    static const char H[] = "0123456789abcdef";
    for (int i = 0; i < len; i++) { out[2*i] = H[in[i] >> 4]; out[2*i+1] = H[in[i] & 15]; }
    out[2*len] = 0;
}
/* "0x" + EIP-55 mixed-case of a 20-byte address; out >= 43 bytes */
static inline void eip55_encode(const uint8_t addr[20], char *out) {
    char lower[41]; uint8_t h[32];
    hex_encode(addr, 20, lower); keccak256((const uint8_t *)lower, 40, h);
    out[0] = '0'; out[1] = 'x';
    for (int i = 0; i < 40; i++) {
        int nib = (i & 1) ? (h[i >> 1] & 15) : (h[i >> 1] >> 4);
        char c = lower[i];
        out[2 + i] = (c >= 'a' && nib >= 8) ? (char)(c - 32) : c;
    }
    out[42] = 0;
}
static inline int wif_encode(const uint8_t key[32], int compressed, char *out) {
    uint8_t p[34]; p[0] = 0x80; memcpy(p + 1, key, 32);
    if (compressed) { p[33] = 0x01; return base58check_encode(p, 34, out); }
    return base58check_encode(p, 33, out);
}
#endif
