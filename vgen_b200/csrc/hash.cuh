// Register-resident single-message hashes for the scan kernels: SHA-256 (33/65/22/21/32-byte
// messages), RIPEMD-160 (32-byte message), Keccak-256 (64- and 40-byte messages).
//
// Replaces src/shaders/sha256.wgsl:43-170 and src/shaders/ripemd160.wgsl:10-100 (one-block
// versions with the padding baked in) and adds what the reference left on the host: the second
// hash160 of P2SH-P2WPKH (src/gpu.rs:1051-1060), the uncompressed-key hash
// (src/address.rs:121-124) and Keccak-256 (src/address.rs:95-113,176-198).
// All loops are fully unrolled; message words that are padding constants fold at compile time.
#pragma once
#include <stdint.h>
#include "fe.cuh"

#ifdef VG_HOST_EMU
static inline uint32_t vg_rotr(uint32_t x, int n) { return (x >> n) | (x << (32 - n)); }
static inline uint32_t vg_rotl(uint32_t x, int n) { return (x << n) | (x >> (32 - n)); }
static inline uint32_t vg_bswap(uint32_t x) { return __builtin_bswap32(x); }
static inline uint32_t vg_shf_l(uint32_t lo, uint32_t hi, int n) { return (hi << n) | (lo >> (32 - n)); }  // 0<n<32
#define VG_CONST static const
// pipe-steering primitives (device: forced onto the multiplier pipe; host emulation: plain arithmetic)
static inline uint32_t vg_addf(uint32_t a, uint32_t b) { return a + b; }
static inline uint32_t vg_rotl_f(uint32_t x, int n) { return vg_rotl(x, n); }
static inline uint32_t vg_shr_f(uint32_t x, int n) { return x >> n; }
#else
__device__ __forceinline__ uint32_t vg_rotr(uint32_t x, int n) { return __funnelshift_r(x, x, n); }
__device__ __forceinline__ uint32_t vg_rotl(uint32_t x, int n) { return __funnelshift_l(x, x, n); }
__device__ __forceinline__ uint32_t vg_bswap(uint32_t x) { return __byte_perm(x, 0, 0x0123); }
__device__ __forceinline__ uint32_t vg_shf_l(uint32_t lo, uint32_t hi, int n) { return __funnelshift_l(lo, hi, n); }
#define VG_CONST __device__ __constant__ const
// ---- pipe steering -------------------------------------------------------------------------------------------
// On sm_100 the add/logic/shift instructions (IADD3, LOP3, SHF, LEA) issue on the 64-lane ALU pipe and IMAD /
// IMAD.WIDE on the 64-lane multiplier pipe. SHA-256 + RIPEMD-160 are ~95 % ALU-class work, so the ALU pipe binds
// (ncu r01_a: ALU 74 % busy, multiplier pipe 32 %). These helpers move selected adds, rotations and shifts onto the
// multiplier pipe. The multiplier 1 is read from constant memory so ptxas cannot fold the multiply back into an add.
// MEASURED (B200, r01, tools/variants.sh): every steering variant is SLOWER than leaving pipe selection to ptxas
// (5.12-5.37 vs 5.82 Gkeys/s): peak_bench shows the two pipes do not add up on this chip (pure ALU 19.3, pure IMAD 16.8,
// 1:1 mix only 25.6 T inst/s), so trading one ALU op for 1-2 IMADs loses. All three switches therefore default to 0.
#ifndef VG_STEER_ADD
#define VG_STEER_ADD 0
#endif
#ifndef VG_STEER_ROT
#define VG_STEER_ROT 0
#endif
#ifndef VG_STEER_SHR
#define VG_STEER_SHR 0
#endif
__device__ __constant__ uint32_t vg_one = 1;
// a + b as IMAD (a * 1 + b). Callers use plain + where an operand pair is a compile-time constant.
__device__ __forceinline__ uint32_t vg_addf(uint32_t a, uint32_t b) {
    if (!VG_STEER_ADD) return a + b;
    uint32_t r;
    asm("mad.lo.u32 %0,%1,%2,%3;" : "=r"(r) : "r"(a), "r"(vg_one), "r"(b));
    return r;
}
// rotl(x, n) = lo + hi of the 64-bit product x * 2^n (IMAD.WIDE + IMAD): two multiplier-pipe ops, no ALU op
__device__ __forceinline__ uint32_t vg_rotl_f(uint32_t x, int n) {
    if (!VG_STEER_ROT) return __funnelshift_l(x, x, n);
    uint32_t lo, hi, r;
    asm("{ .reg .u64 p; mul.wide.u32 p,%2,%3; mov.b64 {%0,%1},p; }" : "=r"(lo), "=r"(hi) : "r"(x), "r"(1u << n));
    asm("mad.lo.u32 %0,%1,%2,%3;" : "=r"(r) : "r"(lo), "r"(vg_one), "r"(hi));
    return r;
}
// x >> n as the high half of x * 2^(32-n)
__device__ __forceinline__ uint32_t vg_shr_f(uint32_t x, int n) {
    if (!VG_STEER_SHR) return x >> n;
    uint32_t r;
    asm("mul.hi.u32 %0,%1,%2;" : "=r"(r) : "r"(x), "r"(1u << (32 - n)));
    return r;
}
#endif

// ---- N independent hash instances advanced in lock-step ------------------------------------------------------------
// The hashes below are written over a word type T: uint32_t, or vw<N> = N independent 32-bit lanes held by ONE thread.
// The scan kernel hashes the +i and -i neighbours of a centre as vw<2>: two independent dependency chains in the same
// instruction stream, which is what fills the issue slots a single SHA-256/RIPEMD-160 chain leaves empty (ncu r01_a:
// 49 % issue utilisation, top stall = fixed-latency dependency wait).
template <int N> struct vw { uint32_t v[N]; };
#define VW_BIN(OP) \
    template <int N> VG_DEV vw<N> operator OP(const vw<N> &a, const vw<N> &b) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = a.v[i] OP b.v[i]; return r; } \
    template <int N> VG_DEV vw<N> operator OP(const vw<N> &a, uint32_t b) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = a.v[i] OP b; return r; } \
    template <int N> VG_DEV vw<N> operator OP(uint32_t a, const vw<N> &b) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = a OP b.v[i]; return r; }
VW_BIN(+) VW_BIN(^) VW_BIN(&) VW_BIN(|)
#undef VW_BIN
template <int N> VG_DEV vw<N> operator~(const vw<N> &a) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = ~a.v[i]; return r; }
template <int N> VG_DEV vw<N> operator<<(const vw<N> &a, int n) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = a.v[i] << n; return r; }
template <int N> VG_DEV vw<N> operator>>(const vw<N> &a, int n) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = a.v[i] >> n; return r; }
#define VW_UN(F) template <int N> VG_DEV vw<N> F(const vw<N> &a, int n) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = F(a.v[i], n); return r; }
VW_UN(vg_rotr) VW_UN(vg_rotl) VW_UN(vg_rotl_f) VW_UN(vg_shr_f)
#undef VW_UN
template <int N> VG_DEV vw<N> vg_bswap(const vw<N> &a) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = vg_bswap(a.v[i]); return r; }
template <int N> VG_DEV vw<N> vg_shf_l(const vw<N> &lo, const vw<N> &hi, int n) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = vg_shf_l(lo.v[i], hi.v[i], n); return r; }
template <int N> VG_DEV vw<N> vg_addf(const vw<N> &a, const vw<N> &b) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = vg_addf(a.v[i], b.v[i]); return r; }
template <int N> VG_DEV vw<N> vg_addf(const vw<N> &a, uint32_t b) { vw<N> r; VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = vg_addf(a.v[i], b); return r; }
VG_DEV void vg_splat(uint32_t &r, uint32_t c) { r = c; }
template <int N> VG_DEV void vg_splat(vw<N> &r, uint32_t c) { VG_UNROLL for (int i = 0; i < N; i++) r.v[i] = c; }

VG_CONST uint32_t VG_SHA_K[64] = {
    0x428a2f98,0x71374491,0xb5c0fbcf,0xe9b5dba5,0x3956c25b,0x59f111f1,0x923f82a4,0xab1c5ed5,
    0xd807aa98,0x12835b01,0x243185be,0x550c7dc3,0x72be5d74,0x80deb1fe,0x9bdc06a7,0xc19bf174,
    0xe49b69c1,0xefbe4786,0x0fc19dc6,0x240ca1cc,0x2de92c6f,0x4a7484aa,0x5cb0a9dc,0x76f988da,
    0x983e5152,0xa831c66d,0xb00327c8,0xbf597fc7,0xc6e00bf3,0xd5a79147,0x06ca6351,0x14292967,
    0x27b70a85,0x2e1b2138,0x4d2c6dfc,0x53380d13,0x650a7354,0x766a0abb,0x81c2c92e,0x92722c85,
    0xa2bfe8a1,0xa81a664b,0xc24b8b70,0xc76c51a3,0xd192e819,0xd6990624,0xf40e3585,0x106aa070,
    0x19a4c116,0x1e376c08,0x2748774c,0x34b0bcb5,0x391c0cb3,0x4ed8aa4a,0x5b9cca4f,0x682e6ff3,
    0x748f82ee,0x78a5636f,0x84c87814,0x8cc70208,0x90befffa,0xa4506ceb,0xbef9a3f7,0xc67178f2};

#define VG_SHA_IV {0x6a09e667u,0xbb67ae85u,0x3c6ef372u,0xa54ff53au,0x510e527fu,0x9b05688cu,0x1f83d9abu,0x5be0cd19u}
template <class T> VG_DEV void sha256_iv(T st[8]) {
    const uint32_t iv[8] = VG_SHA_IV;
    VG_UNROLL for (int i = 0; i < 8; i++) vg_splat(st[i], iv[i]);
}

// How many of the six rotations per SHA-256 round go to the multiplier pipe (0..2); tuned on the B200, see DESIGN.md
#ifndef VG_SHA_ROT_FMA
#define VG_SHA_ROT_FMA 0
#endif
// one compression; w[] is consumed as the rolling schedule.
// Pipe split per round: S1, ch, S0, maj, t = p + S1 + ch, a = t + S0 + maj on the ALU pipe (12 ops);
// K + w, p = h + (K + w), e = d + t on the multiplier pipe (3 ops). Schedule: 2 x (2 rotations + LOP3) + one 3-input add
// on the ALU pipe, the two plain shifts and w[i-16] + w[i-7] on the multiplier pipe.
// CMASK: bit i set = message word i is a compile-time constant (padding), so K + w folds and must not be forced
// through an IMAD.
// Hooks. The hashes below are ~2,300 add/logic/shift instructions in a row: pure ALU-pipe work. A caller may pass a hook object
// whose at(k) is called at fixed points between rounds (k = 0..7 after every 8th SHA-256 round, k = 8..17 after every 16th
// RIPEMD-160 step); the walk kernel uses it to issue the NEXT point's field multiplications (multiplier-pipe work) in between
// (scan_core.cuh VG_PIPELINE), so that one warp feeds both pipes instead of taking turns. k is a compile-time constant after
// unrolling, so an at() that tests k folds to nothing where it does nothing.
struct vg_no_hook { VG_DEVM void at(int) const {} };
template <uint32_t CMASK, class T, class Hook>
VG_DEV void sha256_compress(T st[8], T w[16], Hook &hk) {
    T a = st[0], b = st[1], c = st[2], d = st[3], e = st[4], f = st[5], g = st[6], h = st[7];
    VG_UNROLL for (int i = 0; i < 64; i++) {
        if (i >= 16) {
            T w15 = w[(i - 15) & 15], w2 = w[(i - 2) & 15];
            T s0 = vg_rotr(w15, 7) ^ vg_rotr(w15, 18) ^ vg_shr_f(w15, 3);
            T s1 = vg_rotr(w2, 17) ^ vg_rotr(w2, 19) ^ vg_shr_f(w2, 10);
            w[i & 15] = vg_addf(w[i & 15], w[(i - 7) & 15]) + s0 + s1;
        }
        T S1 = vg_rotr(e, 6) ^ vg_rotr(e, 11) ^ (VG_SHA_ROT_FMA >= 1 ? vg_rotl_f(e, 7) : vg_rotr(e, 25));
        T ch = (e & f) ^ (~e & g);
        T kw = (i < 16 && ((CMASK >> i) & 1u)) ? w[i & 15] + VG_SHA_K[i] : vg_addf(w[i & 15], VG_SHA_K[i]);
        T p = vg_addf(h, kw);
        T t1 = p + S1 + ch;
        T S0 = vg_rotr(a, 2) ^ vg_rotr(a, 13) ^ (VG_SHA_ROT_FMA >= 2 ? vg_rotl_f(a, 10) : vg_rotr(a, 22));
        T mj = (a & b) | (c & (a | b));
        h = g; g = f; f = e; e = vg_addf(d, t1); d = c; c = b; b = a; a = t1 + S0 + mj;
        if ((i & 7) == 7) hk.at(i >> 3);
    }
    st[0] = vg_addf(st[0], a); st[1] = vg_addf(st[1], b); st[2] = vg_addf(st[2], c); st[3] = vg_addf(st[3], d);
    st[4] = vg_addf(st[4], e); st[5] = vg_addf(st[5], f); st[6] = vg_addf(st[6], g); st[7] = vg_addf(st[7], h);
}
template <uint32_t CMASK, class T>
VG_DEV void sha256_compress(T st[8], T w[16]) { vg_no_hook nh; sha256_compress<CMASK, T, vg_no_hook>(st, w, nh); }
// SHA-256(prefix || X) for a 33-byte compressed key; xv = the 8 little-endian limbs of a canonical field element
template <class T, class Hook>
VG_DEV void sha256_pub33(T out[8], const T &prefix, const T xv[8], Hook &hk) {
    T w[16];
    w[0] = (prefix << 24) | (xv[7] >> 8);
    VG_UNROLL for (int k = 1; k < 8; k++) w[k] = vg_shf_l(xv[7 - k], xv[8 - k], 24);
    w[8] = (xv[0] << 24) | 0x00800000u;
    VG_UNROLL for (int k = 9; k < 15; k++) vg_splat(w[k], 0u);
    vg_splat(w[15], 33 * 8);
    sha256_iv(out);
    sha256_compress<0xFE00u, T, Hook>(out, w, hk);
}
template <class T>
VG_DEV void sha256_pub33(T out[8], const T &prefix, const T xv[8]) { vg_no_hook nh; sha256_pub33<T, vg_no_hook>(out, prefix, xv, nh); }
// SHA-256(04 || X || Y) for a 65-byte uncompressed key (two blocks)
template <class T>
VG_DEV void sha256_pub65(T out[8], const T xv[8], const T yv[8]) {
    T w[16];
    w[0] = (xv[7] >> 8) | (4u << 24);
    VG_UNROLL for (int k = 1; k < 8; k++) w[k] = vg_shf_l(xv[7 - k], xv[8 - k], 24);
    w[8] = (xv[0] << 24) | (yv[7] >> 8);
    VG_UNROLL for (int k = 1; k < 8; k++) w[8 + k] = vg_shf_l(yv[7 - k], yv[8 - k], 24);
    sha256_iv(out);
    sha256_compress<0u>(out, w);
    w[0] = (yv[0] << 24) | 0x00800000u;
    VG_UNROLL for (int k = 1; k < 15; k++) vg_splat(w[k], 0u);
    vg_splat(w[15], 65 * 8);
    sha256_compress<0xFFFEu>(out, w);
}
// SHA-256 of a message of NBYTES (<= 55) given as big-endian words (unused tail bytes must be 0)
template <int NBYTES, class T>
VG_DEV void sha256_short(T out[8], const T *m) {
    const int nbytes = NBYTES;
    T w[16];
    VG_UNROLL for (int k = 0; k < 15; k++) { if (k * 4 < nbytes + 1) w[k] = m[k]; else vg_splat(w[k], 0u); }
    w[nbytes >> 2] = w[nbytes >> 2] | (0x80u << (24 - 8 * (nbytes & 3)));
    vg_splat(w[15], (uint32_t)nbytes * 8);
    sha256_iv(out);
    sha256_compress<(0xFFFFu << (NBYTES / 4 + 1)) & 0xFFFFu>(out, w);
}

// ---- RIPEMD-160 of a 32-byte message given as 8 big-endian SHA words ----
#define VG_RMD_F0(x, y, z) ((x) ^ (y) ^ (z))
#define VG_RMD_F1(x, y, z) (((x) & (y)) | (~(x) & (z)))
#define VG_RMD_F2(x, y, z) (((x) | ~(y)) ^ (z))
#define VG_RMD_F3(x, y, z) (((x) & (z)) | ((y) & ~(z)))
#define VG_RMD_F4(x, y, z) ((x) ^ ((y) | ~(z)))
// ALU pipe: F (one LOP3), a + F + (x + k) (one IADD3), rotate-and-add e (one LEA.HI); multiplier pipe: x + k, rotl(c, 10)
#define VG_RMD_STEP(F, a, b, c, d, e, xw, s) \
    a = vg_rotl(a + F(b, c, d) + (xw), s) + e; c = vg_rotl_f(c, 10);

template <class T, class Hook>
VG_DEV void ripemd160_32(T out[5], const T sha[8], Hook &hk) {
    T x[16];
    VG_UNROLL for (int i = 0; i < 8; i++) x[i] = vg_bswap(sha[i]);
    const uint32_t pad[8] = {0x80u, 0, 0, 0, 0, 0, 256u, 0};
    VG_UNROLL for (int i = 0; i < 8; i++) vg_splat(x[8 + i], pad[i]);
    const uint32_t h0 = 0x67452301u, h1 = 0xEFCDAB89u, h2 = 0x98BADCFEu, h3 = 0x10325476u, h4 = 0xC3D2E1F0u;
    T a, b, c, d, e, A, B, C, D, E;
    vg_splat(a, h0); vg_splat(b, h1); vg_splat(c, h2); vg_splat(d, h3); vg_splat(e, h4);
    A = a; B = b; C = c; D = d; E = e;
// x[8..15] are padding constants: x + k folds at compile time there
#define L(F, a, b, c, d, e, r, s, k) VG_RMD_STEP(F, a, b, c, d, e, ((r) >= 8 || (k) == 0u ? x[r] + (k) : vg_addf(x[r], k)), s)
    // left line
    L(VG_RMD_F0,a,b,c,d,e,0,11,0u) L(VG_RMD_F0,e,a,b,c,d,1,14,0u) L(VG_RMD_F0,d,e,a,b,c,2,15,0u) L(VG_RMD_F0,c,d,e,a,b,3,12,0u)
    L(VG_RMD_F0,b,c,d,e,a,4,5,0u) L(VG_RMD_F0,a,b,c,d,e,5,8,0u) L(VG_RMD_F0,e,a,b,c,d,6,7,0u) L(VG_RMD_F0,d,e,a,b,c,7,9,0u)
    L(VG_RMD_F0,c,d,e,a,b,8,11,0u) L(VG_RMD_F0,b,c,d,e,a,9,13,0u) L(VG_RMD_F0,a,b,c,d,e,10,14,0u) L(VG_RMD_F0,e,a,b,c,d,11,15,0u)
    L(VG_RMD_F0,d,e,a,b,c,12,6,0u) L(VG_RMD_F0,c,d,e,a,b,13,7,0u) L(VG_RMD_F0,b,c,d,e,a,14,9,0u) L(VG_RMD_F0,a,b,c,d,e,15,8,0u)
    hk.at(8);
#define K1 0x5A827999u
    L(VG_RMD_F1,e,a,b,c,d,7,7,K1) L(VG_RMD_F1,d,e,a,b,c,4,6,K1) L(VG_RMD_F1,c,d,e,a,b,13,8,K1) L(VG_RMD_F1,b,c,d,e,a,1,13,K1)
    L(VG_RMD_F1,a,b,c,d,e,10,11,K1) L(VG_RMD_F1,e,a,b,c,d,6,9,K1) L(VG_RMD_F1,d,e,a,b,c,15,7,K1) L(VG_RMD_F1,c,d,e,a,b,3,15,K1)
    L(VG_RMD_F1,b,c,d,e,a,12,7,K1) L(VG_RMD_F1,a,b,c,d,e,0,12,K1) L(VG_RMD_F1,e,a,b,c,d,9,15,K1) L(VG_RMD_F1,d,e,a,b,c,5,9,K1)
    L(VG_RMD_F1,c,d,e,a,b,2,11,K1) L(VG_RMD_F1,b,c,d,e,a,14,7,K1) L(VG_RMD_F1,a,b,c,d,e,11,13,K1) L(VG_RMD_F1,e,a,b,c,d,8,12,K1)
    hk.at(9);
#define K2 0x6ED9EBA1u
    L(VG_RMD_F2,d,e,a,b,c,3,11,K2) L(VG_RMD_F2,c,d,e,a,b,10,13,K2) L(VG_RMD_F2,b,c,d,e,a,14,6,K2) L(VG_RMD_F2,a,b,c,d,e,4,7,K2)
    L(VG_RMD_F2,e,a,b,c,d,9,14,K2) L(VG_RMD_F2,d,e,a,b,c,15,9,K2) L(VG_RMD_F2,c,d,e,a,b,8,13,K2) L(VG_RMD_F2,b,c,d,e,a,1,15,K2)
    L(VG_RMD_F2,a,b,c,d,e,2,14,K2) L(VG_RMD_F2,e,a,b,c,d,7,8,K2) L(VG_RMD_F2,d,e,a,b,c,0,13,K2) L(VG_RMD_F2,c,d,e,a,b,6,6,K2)
    L(VG_RMD_F2,b,c,d,e,a,13,5,K2) L(VG_RMD_F2,a,b,c,d,e,11,12,K2) L(VG_RMD_F2,e,a,b,c,d,5,7,K2) L(VG_RMD_F2,d,e,a,b,c,12,5,K2)
    hk.at(10);
#define K3 0x8F1BBCDCu
    L(VG_RMD_F3,c,d,e,a,b,1,11,K3) L(VG_RMD_F3,b,c,d,e,a,9,12,K3) L(VG_RMD_F3,a,b,c,d,e,11,14,K3) L(VG_RMD_F3,e,a,b,c,d,10,15,K3)
    L(VG_RMD_F3,d,e,a,b,c,0,14,K3) L(VG_RMD_F3,c,d,e,a,b,8,15,K3) L(VG_RMD_F3,b,c,d,e,a,12,9,K3) L(VG_RMD_F3,a,b,c,d,e,4,8,K3)
    L(VG_RMD_F3,e,a,b,c,d,13,9,K3) L(VG_RMD_F3,d,e,a,b,c,3,14,K3) L(VG_RMD_F3,c,d,e,a,b,7,5,K3) L(VG_RMD_F3,b,c,d,e,a,15,6,K3)
    L(VG_RMD_F3,a,b,c,d,e,14,8,K3) L(VG_RMD_F3,e,a,b,c,d,5,6,K3) L(VG_RMD_F3,d,e,a,b,c,6,5,K3) L(VG_RMD_F3,c,d,e,a,b,2,12,K3)
    hk.at(11);
#define K4 0xA953FD4Eu
    L(VG_RMD_F4,b,c,d,e,a,4,9,K4) L(VG_RMD_F4,a,b,c,d,e,0,15,K4) L(VG_RMD_F4,e,a,b,c,d,5,5,K4) L(VG_RMD_F4,d,e,a,b,c,9,11,K4)
    L(VG_RMD_F4,c,d,e,a,b,7,6,K4) L(VG_RMD_F4,b,c,d,e,a,12,8,K4) L(VG_RMD_F4,a,b,c,d,e,2,13,K4) L(VG_RMD_F4,e,a,b,c,d,10,12,K4)
    L(VG_RMD_F4,d,e,a,b,c,14,5,K4) L(VG_RMD_F4,c,d,e,a,b,1,12,K4) L(VG_RMD_F4,b,c,d,e,a,3,13,K4) L(VG_RMD_F4,a,b,c,d,e,8,14,K4)
    L(VG_RMD_F4,e,a,b,c,d,11,11,K4) L(VG_RMD_F4,d,e,a,b,c,6,8,K4) L(VG_RMD_F4,c,d,e,a,b,15,5,K4) L(VG_RMD_F4,b,c,d,e,a,13,6,K4)
    hk.at(12);
    // right line
#define J0 0x50A28BE6u
    L(VG_RMD_F4,A,B,C,D,E,5,8,J0) L(VG_RMD_F4,E,A,B,C,D,14,9,J0) L(VG_RMD_F4,D,E,A,B,C,7,9,J0) L(VG_RMD_F4,C,D,E,A,B,0,11,J0)
    L(VG_RMD_F4,B,C,D,E,A,9,13,J0) L(VG_RMD_F4,A,B,C,D,E,2,15,J0) L(VG_RMD_F4,E,A,B,C,D,11,15,J0) L(VG_RMD_F4,D,E,A,B,C,4,5,J0)
    L(VG_RMD_F4,C,D,E,A,B,13,7,J0) L(VG_RMD_F4,B,C,D,E,A,6,7,J0) L(VG_RMD_F4,A,B,C,D,E,15,8,J0) L(VG_RMD_F4,E,A,B,C,D,8,11,J0)
    L(VG_RMD_F4,D,E,A,B,C,1,14,J0) L(VG_RMD_F4,C,D,E,A,B,10,14,J0) L(VG_RMD_F4,B,C,D,E,A,3,12,J0) L(VG_RMD_F4,A,B,C,D,E,12,6,J0)
    hk.at(13);
#define J1 0x5C4DD124u
    L(VG_RMD_F3,E,A,B,C,D,6,9,J1) L(VG_RMD_F3,D,E,A,B,C,11,13,J1) L(VG_RMD_F3,C,D,E,A,B,3,15,J1) L(VG_RMD_F3,B,C,D,E,A,7,7,J1)
    L(VG_RMD_F3,A,B,C,D,E,0,12,J1) L(VG_RMD_F3,E,A,B,C,D,13,8,J1) L(VG_RMD_F3,D,E,A,B,C,5,9,J1) L(VG_RMD_F3,C,D,E,A,B,10,11,J1)
    L(VG_RMD_F3,B,C,D,E,A,14,7,J1) L(VG_RMD_F3,A,B,C,D,E,15,7,J1) L(VG_RMD_F3,E,A,B,C,D,8,12,J1) L(VG_RMD_F3,D,E,A,B,C,12,7,J1)
    L(VG_RMD_F3,C,D,E,A,B,4,6,J1) L(VG_RMD_F3,B,C,D,E,A,9,15,J1) L(VG_RMD_F3,A,B,C,D,E,1,13,J1) L(VG_RMD_F3,E,A,B,C,D,2,11,J1)
    hk.at(14);
#define J2 0x6D703EF3u
    L(VG_RMD_F2,D,E,A,B,C,15,9,J2) L(VG_RMD_F2,C,D,E,A,B,5,7,J2) L(VG_RMD_F2,B,C,D,E,A,1,15,J2) L(VG_RMD_F2,A,B,C,D,E,3,11,J2)
    L(VG_RMD_F2,E,A,B,C,D,7,8,J2) L(VG_RMD_F2,D,E,A,B,C,14,6,J2) L(VG_RMD_F2,C,D,E,A,B,6,6,J2) L(VG_RMD_F2,B,C,D,E,A,9,14,J2)
    L(VG_RMD_F2,A,B,C,D,E,11,12,J2) L(VG_RMD_F2,E,A,B,C,D,8,13,J2) L(VG_RMD_F2,D,E,A,B,C,12,5,J2) L(VG_RMD_F2,C,D,E,A,B,2,14,J2)
    L(VG_RMD_F2,B,C,D,E,A,10,13,J2) L(VG_RMD_F2,A,B,C,D,E,0,13,J2) L(VG_RMD_F2,E,A,B,C,D,4,7,J2) L(VG_RMD_F2,D,E,A,B,C,13,5,J2)
    hk.at(15);
#define J3 0x7A6D76E9u
    L(VG_RMD_F1,C,D,E,A,B,8,15,J3) L(VG_RMD_F1,B,C,D,E,A,6,5,J3) L(VG_RMD_F1,A,B,C,D,E,4,8,J3) L(VG_RMD_F1,E,A,B,C,D,1,11,J3)
    L(VG_RMD_F1,D,E,A,B,C,3,14,J3) L(VG_RMD_F1,C,D,E,A,B,11,14,J3) L(VG_RMD_F1,B,C,D,E,A,15,6,J3) L(VG_RMD_F1,A,B,C,D,E,0,14,J3)
    L(VG_RMD_F1,E,A,B,C,D,5,6,J3) L(VG_RMD_F1,D,E,A,B,C,12,9,J3) L(VG_RMD_F1,C,D,E,A,B,2,12,J3) L(VG_RMD_F1,B,C,D,E,A,13,9,J3)
    L(VG_RMD_F1,A,B,C,D,E,9,12,J3) L(VG_RMD_F1,E,A,B,C,D,7,5,J3) L(VG_RMD_F1,D,E,A,B,C,10,15,J3) L(VG_RMD_F1,C,D,E,A,B,14,8,J3)
    hk.at(16);
    L(VG_RMD_F0,B,C,D,E,A,12,8,0u) L(VG_RMD_F0,A,B,C,D,E,15,5,0u) L(VG_RMD_F0,E,A,B,C,D,10,12,0u) L(VG_RMD_F0,D,E,A,B,C,4,9,0u)
    L(VG_RMD_F0,C,D,E,A,B,1,12,0u) L(VG_RMD_F0,B,C,D,E,A,5,5,0u) L(VG_RMD_F0,A,B,C,D,E,8,14,0u) L(VG_RMD_F0,E,A,B,C,D,7,6,0u)
    L(VG_RMD_F0,D,E,A,B,C,6,8,0u) L(VG_RMD_F0,C,D,E,A,B,2,13,0u) L(VG_RMD_F0,B,C,D,E,A,13,6,0u) L(VG_RMD_F0,A,B,C,D,E,14,5,0u)
    L(VG_RMD_F0,E,A,B,C,D,0,15,0u) L(VG_RMD_F0,D,E,A,B,C,3,13,0u) L(VG_RMD_F0,C,D,E,A,B,9,11,0u) L(VG_RMD_F0,B,C,D,E,A,11,11,0u)
#undef L
    hk.at(17);
    out[0] = c + D + h1; out[1] = d + E + h2; out[2] = e + A + h3; out[3] = a + B + h4; out[4] = b + C + h0;
}
template <class T>
VG_DEV void ripemd160_32(T out[5], const T sha[8]) { vg_no_hook nh; ripemd160_32<T, vg_no_hook>(out, sha, nh); }
// hash160 words (little-endian: digest byte 4i+j = byte j of out[i])
template <class T, class Hook>
VG_DEV void hash160_pub33(T out[5], const T &prefix, const T xv[8], Hook &hk) {
    T s[8]; sha256_pub33<T, Hook>(s, prefix, xv, hk); ripemd160_32<T, Hook>(out, s, hk);
}
template <class T>
VG_DEV void hash160_pub33(T out[5], const T &prefix, const T xv[8]) { vg_no_hook nh; hash160_pub33<T, vg_no_hook>(out, prefix, xv, nh); }
template <class T>
VG_DEV void hash160_pub65(T out[5], const T xv[8], const T yv[8]) {
    T s[8]; sha256_pub65(s, xv, yv); ripemd160_32(out, s);
}
// hash160(00 14 || h) : the P2SH-P2WPKH redeem-script hash (src/address.rs:130-135)
template <class T>
VG_DEV void hash160_p2sh_script(T out[5], const T h[5]) {
    T B[5], mm[15], s[8];
    VG_UNROLL for (int i = 0; i < 5; i++) B[i] = vg_bswap(h[i]);
    mm[0] = (B[0] >> 16) | 0x00140000u;
    VG_UNROLL for (int i = 1; i < 5; i++) mm[i] = (B[i - 1] << 16) | (B[i] >> 16);
    mm[5] = B[4] << 16;
    VG_UNROLL for (int i = 6; i < 15; i++) vg_splat(mm[i], 0u);
    sha256_short<22>(s, mm);
    ripemd160_32(out, s);
}

// ---- Keccak-256 (0x01 padding), state as 25 x 64-bit lanes ----
VG_CONST uint64_t VG_KECCAK_RC[24] = {
    0x0000000000000001ULL,0x0000000000008082ULL,0x800000000000808aULL,0x8000000080008000ULL,
    0x000000000000808bULL,0x0000000080000001ULL,0x8000000080008081ULL,0x8000000000008009ULL,
    0x000000000000008aULL,0x0000000000000088ULL,0x0000000080008009ULL,0x000000008000000aULL,
    0x000000008000808bULL,0x800000000000008bULL,0x8000000000008089ULL,0x8000000000008003ULL,
    0x8000000000008002ULL,0x8000000000000080ULL,0x000000000000800aULL,0x800000008000000aULL,
    0x8000000080008081ULL,0x8000000000008080ULL,0x0000000080000001ULL,0x8000000080008008ULL};

// 64-bit lanes are kept as two 32-bit halves (L = low, H = high) so that every rotation half is exactly one funnel
// shift (SHF) and every xor / chi one LOP3. Written on uint64_t, nvcc lowers part of the rotations to shift + shift + or
// and rotl-by-1 to add-with-carry: 248 instructions per round against 180 here (ncu: 7,021 -> see DESIGN.md).
template <int N> VG_DEV void vg_rotl64_halves(uint32_t &ol, uint32_t &oh, uint32_t l, uint32_t h) {
    if (N == 0) { ol = l; oh = h; }
    else if (N < 32) { oh = vg_shf_l(l, h, N & 31); ol = vg_shf_l(h, l, N & 31); }
    else if (N == 32) { ol = h; oh = l; }
    else { oh = vg_shf_l(h, l, (N - 32) & 31); ol = vg_shf_l(l, h, (N - 32) & 31); }
}
VG_DEV uint32_t vg_xor3(uint32_t a, uint32_t b, uint32_t c) {
#ifdef VG_HOST_EMU
    return a ^ b ^ c;
#else
    uint32_t r;
    asm("lop3.b32 %0,%1,%2,%3,0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
#endif
}
// b[DST] = rotl64(s[SRC] ^ d[SRC % 5], N), with d[x] = c[x-1] ^ rotl64(c[x+1], 1) folded into a 3-input xor
#define VG_K_RHO(DST, SRC, N) vg_rotl64_halves<N>(bl[DST], bh[DST], vg_xor3(sl[SRC], ca_l[(SRC) % 5], r1_l[(SRC) % 5]), vg_xor3(sh[SRC], ca_h[(SRC) % 5], r1_h[(SRC) % 5]))

// VG_KECCAK_UNROLL: 1 = fully unrolled rounds (round-constant loads fold, the zero lanes of the first round and the
// unused lanes of the last one are eliminated), 0 = rolled loop (smaller code). Measured on B200 (64-bit version):
// unrolled is 9 % slower (instruction cache).
// VG_KECCAK_PEEL (rolled loop only): the first and the last round are separate copies of the round body. In the first,
// 17 of the 25 input lanes are zero or padding constants and fold away; in the last, only row 0 of the state is read by
// the callers (lanes 0..3), so 20 of the 25 rho/pi lanes and 4 of the 5 chi rows are dead code. After keccak_f1600 only
// lanes 0..4 are meaningful.
#ifndef VG_KECCAK_UNROLL
#define VG_KECCAK_UNROLL 0
#endif
#ifndef VG_KECCAK_PEEL
#define VG_KECCAK_PEEL 1
#endif
VG_DEV void keccak_round(uint32_t sl[25], uint32_t sh[25], uint32_t rc_l, uint32_t rc_h) {
    uint32_t cl[5], ch[5];
    VG_UNROLL for (int x = 0; x < 5; x++) {
        cl[x] = vg_xor3(vg_xor3(sl[x], sl[x + 5], sl[x + 10]), sl[x + 15], sl[x + 20]);
        ch[x] = vg_xor3(vg_xor3(sh[x], sh[x + 5], sh[x + 10]), sh[x + 15], sh[x + 20]);
    }
    // for column x: ca = c[x-1], r1 = rotl64(c[x+1], 1)
    uint32_t ca_l[5], ca_h[5], r1_l[5], r1_h[5];
    VG_UNROLL for (int x = 0; x < 5; x++) {
        ca_l[x] = cl[(x + 4) % 5]; ca_h[x] = ch[(x + 4) % 5];
        vg_rotl64_halves<1>(r1_l[x], r1_h[x], cl[(x + 1) % 5], ch[(x + 1) % 5]);
    }
    uint32_t bl[25], bh[25];
    VG_K_RHO(0, 0, 0);
    VG_K_RHO(10, 1, 1);   VG_K_RHO(20, 2, 62);  VG_K_RHO(5, 3, 28);   VG_K_RHO(15, 4, 27);
    VG_K_RHO(16, 5, 36);  VG_K_RHO(1, 6, 44);   VG_K_RHO(11, 7, 6);   VG_K_RHO(21, 8, 55);
    VG_K_RHO(6, 9, 20);   VG_K_RHO(7, 10, 3);   VG_K_RHO(17, 11, 10); VG_K_RHO(2, 12, 43);
    VG_K_RHO(12, 13, 25); VG_K_RHO(22, 14, 39); VG_K_RHO(23, 15, 41); VG_K_RHO(8, 16, 45);
    VG_K_RHO(18, 17, 15); VG_K_RHO(3, 18, 21);  VG_K_RHO(13, 19, 8);  VG_K_RHO(14, 20, 18);
    VG_K_RHO(24, 21, 2);  VG_K_RHO(9, 22, 61);  VG_K_RHO(19, 23, 56); VG_K_RHO(4, 24, 14);
    VG_UNROLL for (int j = 0; j < 25; j += 5) {
        VG_UNROLL for (int x = 0; x < 5; x++) {
            sl[j + x] = bl[j + x] ^ (~bl[j + (x + 1) % 5] & bl[j + (x + 2) % 5]);
            sh[j + x] = bh[j + x] ^ (~bh[j + (x + 1) % 5] & bh[j + (x + 2) % 5]);
        }
    }
    sl[0] ^= rc_l;
    sh[0] ^= rc_h;
}
VG_DEV void keccak_f1600(uint32_t sl[25], uint32_t sh[25]) {
#if VG_KECCAK_PEEL && !VG_KECCAK_UNROLL
    keccak_round(sl, sh, 0x00000001u, 0u);                                  // round 0
  #ifndef VG_HOST_EMU
    #pragma unroll 1
  #endif
    for (int r = 1; r < 23; r++) keccak_round(sl, sh, (uint32_t)VG_KECCAK_RC[r], (uint32_t)(VG_KECCAK_RC[r] >> 32));
    keccak_round(sl, sh, 0x80008008u, 0x80000000u);                         // round 23
#else
#if VG_KECCAK_UNROLL && !defined(VG_HOST_EMU)
    #pragma unroll
#else
  #ifndef VG_HOST_EMU
    #pragma unroll 1
  #endif
#endif
    for (int r = 0; r < 24; r++) keccak_round(sl, sh, (uint32_t)VG_KECCAK_RC[r], (uint32_t)(VG_KECCAK_RC[r] >> 32));
#endif
}
// Ethereum address words: Keccak-256(X || Y)[12..32] as 5 words, digest byte 4i+j = byte j of out[i]
VG_DEV void keccak_eth_address(uint32_t out[5], const fe &x, const fe &y) {
    uint32_t sl[25], sh[25];
    VG_UNROLL for (int i = 0; i < 4; i++) {      // lane i = message bytes 8i..8i+7, little-endian
        sl[i] = vg_bswap(x.v[7 - 2 * i]);     sh[i] = vg_bswap(x.v[6 - 2 * i]);
        sl[4 + i] = vg_bswap(y.v[7 - 2 * i]); sh[4 + i] = vg_bswap(y.v[6 - 2 * i]);
    }
    VG_UNROLL for (int i = 8; i < 25; i++) { sl[i] = 0; sh[i] = 0; }
    sl[8] = 0x01u;
    sh[16] = 0x80000000u;
    keccak_f1600(sl, sh);
    out[0] = sh[1]; out[1] = sl[2]; out[2] = sh[2]; out[3] = sl[3]; out[4] = sh[3];
}
// Keccak-256 of 40 ASCII bytes (the lowercase hex of an address) -> first 20 digest bytes as 5 LE words
VG_DEV void keccak_hex40(uint32_t out[5], const uint32_t m[10]) {   // m: message bytes as little-endian words
    uint32_t sl[25], sh[25];
    VG_UNROLL for (int i = 0; i < 5; i++) { sl[i] = m[2 * i]; sh[i] = m[2 * i + 1]; }
    VG_UNROLL for (int i = 5; i < 25; i++) { sl[i] = 0; sh[i] = 0; }
    sl[5] = 0x01u;
    sh[16] = 0x80000000u;
    keccak_f1600(sl, sh);
    out[0] = sl[0]; out[1] = sh[0]; out[2] = sl[1]; out[3] = sh[1]; out[4] = sl[2];
}
