/* ORACLE (test infrastructure, not product): SHA-256 (FIPS 180-4), RIPEMD-160, Keccak-256
 * (original 0x01 padding, as sha3 0.10.8's Keccak256 used at src/address.rs:100-102,178-180).
 * Generic multi-block byte-oriented versions; the reference reaches them through
 * bitcoin_hashes 0.14.1 (Cargo.lock:182) and sha3 0.10.8 (Cargo.lock:1690). */
#ifndef VGEN_ORACLE_HASHES_H
#define VGEN_ORACLE_HASHES_H
#include <stdint.h>
#include <string.h>
#include <stddef.h>

static inline uint32_t rotr32(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
static inline uint32_t rotl32(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }
static inline uint64_t rotl64(uint64_t x, int n) { return n ? (x << n) | (x >> (64 - n)) : x; }

static const uint32_t SHA_K[64] = {
    0x428a2f98,0x71374491,0xb5c0fbcf,0xe9b5dba5,0x3956c25b,0x59f111f1,0x923f82a4,0xab1c5ed5,
    0xd807aa98,0x12835b01,0x243185be,0x550c7dc3,0x72be5d74,0x80deb1fe,0x9bdc06a7,0xc19bf174,
    0xe49b69c1,0xefbe4786,0x0fc19dc6,0x240ca1cc,0x2de92c6f,0x4a7484aa,0x5cb0a9dc,0x76f988da,
    0x983e5152,0xa831c66d,0xb00327c8,0xbf597fc7,0xc6e00bf3,0xd5a79147,0x06ca6351,0x14292967,
    0x27b70a85,0x2e1b2138,0x4d2c6dfc,0x53380d13,0x650a7354,0x766a0abb,0x81c2c92e,0x92722c85,
    0xa2bfe8a1,0xa81a664b,0xc24b8b70,0xc76c51a3,0xd192e819,0xd6990624,0xf40e3585,0x106aa070,
    0x19a4c116,0x1e376c08,0x2748774c,0x34b0bcb5,0x391c0cb3,0x4ed8aa4a,0x5b9cca4f,0x682e6ff3,
    0x748f82ee,0x78a5636f,0x84c87814,0x8cc70208,0x90befffa,0xa4506ceb,0xbef9a3f7,0xc67178f2};

static inline void sha256_block(uint32_t h[8], const uint8_t blk[64]) {
    uint32_t w[64];
    for (int i = 0; i < 16; i++)
        w[i] = ((uint32_t)blk[4*i] << 24) | ((uint32_t)blk[4*i+1] << 16) | ((uint32_t)blk[4*i+2] << 8) | blk[4*i+3];
    for (int i = 16; i < 64; i++) {
        uint32_t s0 = rotr32(w[i-15], 7) ^ rotr32(w[i-15], 18) ^ (w[i-15] >> 3);
        uint32_t s1 = rotr32(w[i-2], 17) ^ rotr32(w[i-2], 19) ^ (w[i-2] >> 10);
        w[i] = w[i-16] + s0 + w[i-7] + s1;
    }
    uint32_t a=h[0],b=h[1],c=h[2],d=h[3],e=h[4],f=h[5],g=h[6],hh=h[7];
    for (int i = 0; i < 64; i++) {
        uint32_t S1 = rotr32(e,6) ^ rotr32(e,11) ^ rotr32(e,25);
        uint32_t ch = (e & f) ^ (~e & g);
        uint32_t t1 = hh + S1 + ch + SHA_K[i] + w[i];
        uint32_t S0 = rotr32(a,2) ^ rotr32(a,13) ^ rotr32(a,22);
        uint32_t mj = (a & b) ^ (a & c) ^ (b & c);
        uint32_t t2 = S0 + mj;
        hh=g; g=f; f=e; e=d+t1; d=c; c=b; b=a; a=t1+t2;
    }
    h[0]+=a; h[1]+=b; h[2]+=c; h[3]+=d; h[4]+=e; h[5]+=f; h[6]+=g; h[7]+=hh;
}
static inline void sha256_mid(uint32_t h[8], const uint8_t *msg, size_t len, uint64_t total_prior, uint8_t out[32]) {
    /* continue from state h after total_prior bytes (a multiple of 64) */
    size_t off = 0;
    while (len - off >= 64) { sha256_block(h, msg + off); off += 64; }
    uint8_t blk[128]; size_t rem = len - off;
    memset(blk, 0, sizeof blk); memcpy(blk, msg + off, rem); blk[rem] = 0x80;
    size_t n = (rem + 9 <= 64) ? 64 : 128;
    uint64_t bits = (total_prior + len) * 8;
    for (int i = 0; i < 8; i++) blk[n - 1 - i] = (uint8_t)(bits >> (8 * i));
    sha256_block(h, blk); if (n == 128) sha256_block(h, blk + 64);
    for (int i = 0; i < 8; i++) { out[4*i]=h[i]>>24; out[4*i+1]=h[i]>>16; out[4*i+2]=h[i]>>8; out[4*i+3]=h[i]; }
}
static inline void sha256(const uint8_t *msg, size_t len, uint8_t out[32]) {
    uint32_t h[8] = {0x6a09e667,0xbb67ae85,0x3c6ef372,0xa54ff53a,0x510e527f,0x9b05688c,0x1f83d9ab,0x5be0cd19};
    sha256_mid(h, msg, len, 0, out);
}
static inline void sha256d(const uint8_t *msg, size_t len, uint8_t out[32]) {
    uint8_t t[32]; sha256(msg, len, t); sha256(t, 32, out);
}

/* ---- RIPEMD-160 ---- */
static const uint8_t RMD_RL[80] = {
    0,1,2,3,4,5,6,7,8,9,10,11,12,13,14,15, 7,4,13,1,10,6,15,3,12,0,9,5,2,14,11,8,
    3,10,14,4,9,15,8,1,2,7,0,6,13,11,5,12, 1,9,11,10,0,8,12,4,13,3,7,15,14,5,6,2,
    4,0,5,9,7,12,2,10,14,1,3,8,11,6,15,13};
static const uint8_t RMD_RR[80] = {
    5,14,7,0,9,2,11,4,13,6,15,8,1,10,3,12, 6,11,3,7,0,13,5,10,14,15,8,12,4,9,1,2,
    15,5,1,3,7,14,6,9,11,8,12,2,10,0,4,13, 8,6,4,1,3,11,15,0,5,12,2,13,9,7,10,14,
    12,15,10,4,1,5,8,7,6,2,13,14,0,3,9,11};
static const uint8_t RMD_SL[80] = {
    11,14,15,12,5,8,7,9,11,13,14,15,6,7,9,8, 7,6,8,13,11,9,7,15,7,12,15,9,11,7,13,12,
    11,13,6,7,14,9,13,15,14,8,13,6,5,12,7,5, 11,12,14,15,14,15,9,8,9,14,5,6,8,6,5,12,
    9,15,5,11,6,8,13,12,5,12,13,14,11,8,5,6};
static const uint8_t RMD_SR[80] = {
    8,9,9,11,13,15,15,5,7,7,8,11,14,14,12,6, 9,13,15,7,12,8,9,11,7,7,12,7,6,15,13,11,
    9,7,15,11,8,6,6,14,12,13,5,14,13,13,7,5, 15,5,8,11,14,14,6,14,6,9,12,9,12,5,15,8,
    8,5,12,9,12,5,14,6,8,13,6,5,15,13,11,11};
static inline uint32_t rmd_f(int j, uint32_t x, uint32_t y, uint32_t z) {
    switch (j >> 4) {
        case 0: return x ^ y ^ z;
        case 1: return (x & y) | (~x & z);
        case 2: return (x | ~y) ^ z;
        case 3: return (x & z) | (y & ~z);
        default: return x ^ (y | ~z);
    }
}
static inline void ripemd160_block(uint32_t h[5], const uint8_t blk[64]) {
    static const uint32_t KL[5] = {0, 0x5A827999, 0x6ED9EBA1, 0x8F1BBCDC, 0xA953FD4E};
    static const uint32_t KR[5] = {0x50A28BE6, 0x5C4DD124, 0x6D703EF3, 0x7A6D76E9, 0};
    uint32_t x[16];
    for (int i = 0; i < 16; i++)
        x[i] = blk[4*i] | ((uint32_t)blk[4*i+1] << 8) | ((uint32_t)blk[4*i+2] << 16) | ((uint32_t)blk[4*i+3] << 24);
    uint32_t al=h[0],bl=h[1],cl=h[2],dl=h[3],el=h[4], ar=al,br=bl,cr=cl,dr=dl,er=el, t;
    for (int j = 0; j < 80; j++) {
        t = rotl32(al + rmd_f(j, bl, cl, dl) + x[RMD_RL[j]] + KL[j >> 4], RMD_SL[j]) + el;
        al = el; el = dl; dl = rotl32(cl, 10); cl = bl; bl = t;
        t = rotl32(ar + rmd_f(79 - j, br, cr, dr) + x[RMD_RR[j]] + KR[j >> 4], RMD_SR[j]) + er;
        ar = er; er = dr; dr = rotl32(cr, 10); cr = br; br = t;
    }
    t = h[1] + cl + dr; h[1] = h[2] + dl + er; h[2] = h[3] + el + ar;
    h[3] = h[4] + al + br; h[4] = h[0] + bl + cr; h[0] = t;
}
static inline void ripemd160(const uint8_t *msg, size_t len, uint8_t out[20]) {
    uint32_t h[5] = {0x67452301, 0xEFCDAB89, 0x98BADCFE, 0x10325476, 0xC3D2E1F0};
    size_t off = 0;
    while (len - off >= 64) { ripemd160_block(h, msg + off); off += 64; }
    uint8_t blk[128]; size_t rem = len - off;
    memset(blk, 0, sizeof blk); memcpy(blk, msg + off, rem); blk[rem] = 0x80;
    size_t n = (rem + 9 <= 64) ? 64 : 128;
    uint64_t bits = (uint64_t)len * 8;
    for (int i = 0; i < 8; i++) blk[n - 8 + i] = (uint8_t)(bits >> (8 * i));
    ripemd160_block(h, blk); if (n == 128) ripemd160_block(h, blk + 64);
    for (int i = 0; i < 5; i++) { out[4*i]=h[i]; out[4*i+1]=h[i]>>8; out[4*i+2]=h[i]>>16; out[4*i+3]=h[i]>>24; }
}
static inline void hash160(const uint8_t *msg, size_t len, uint8_t out[20]) {
    uint8_t t[32]; sha256(msg, len, t); ripemd160(t, 32, out);
}

/* ---- Keccak-256 ---- */
static inline void keccak_f1600(uint64_t st[25]) {
    static const uint64_t RC[24] = {
        0x0000000000000001ULL,0x0000000000008082ULL,0x800000000000808aULL,0x8000000080008000ULL,
        0x000000000000808bULL,0x0000000080000001ULL,0x8000000080008081ULL,0x8000000000008009ULL,
        0x000000000000008aULL,0x0000000000000088ULL,0x0000000080008009ULL,0x000000008000000aULL,
        0x000000008000808bULL,0x800000000000008bULL,0x8000000000008089ULL,0x8000000000008003ULL,
        0x8000000000008002ULL,0x8000000000000080ULL,0x000000000000800aULL,0x800000008000000aULL,
        0x8000000080008081ULL,0x8000000000008080ULL,0x0000000080000001ULL,0x8000000080008008ULL};
    static const int ROTC[24] = {1,3,6,10,15,21,28,36,45,55,2,14,27,41,56,8,25,43,62,18,39,61,20,44};
    static const int PILN[24] = {10,7,11,17,18,3,5,16,8,21,24,4,15,23,19,13,12,2,20,14,22,9,6,1};
    for (int r = 0; r < 24; r++) {
        uint64_t bc[5], t;
        for (int i = 0; i < 5; i++) bc[i] = st[i] ^ st[i+5] ^ st[i+10] ^ st[i+15] ^ st[i+20];
        for (int i = 0; i < 5; i++) {
            t = bc[(i + 4) % 5] ^ rotl64(bc[(i + 1) % 5], 1);
            for (int j = 0; j < 25; j += 5) st[j + i] ^= t;
        }
        t = st[1];
        for (int i = 0; i < 24; i++) { int j = PILN[i]; uint64_t b = st[j]; st[j] = rotl64(t, ROTC[i]); t = b; }
        for (int j = 0; j < 25; j += 5) {
            for (int i = 0; i < 5; i++) bc[i] = st[j + i];
            for (int i = 0; i < 5; i++) st[j + i] ^= (~bc[(i + 1) % 5]) & bc[(i + 2) % 5];
        }
        st[0] ^= RC[r];
    }
}
static inline void keccak256(const uint8_t *msg, size_t len, uint8_t out[32]) {
    uint64_t st[25]; memset(st, 0, sizeof st);
    const size_t rate = 136; uint8_t blk[136];
    while (len >= rate) {
        for (size_t i = 0; i < rate / 8; i++) { uint64_t v; memcpy(&v, msg + 8*i, 8); st[i] ^= v; }
        keccak_f1600(st); msg += rate; len -= rate;
    }
    memset(blk, 0, rate); memcpy(blk, msg, len); blk[len] ^= 0x01; blk[rate - 1] ^= 0x80;
    for (size_t i = 0; i < rate / 8; i++) { uint64_t v; memcpy(&v, blk + 8*i, 8); st[i] ^= v; }
    keccak_f1600(st);
    memcpy(out, st, 32);
}
#endif
