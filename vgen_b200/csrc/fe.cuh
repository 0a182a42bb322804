// secp256k1 base-field arithmetic for sm_100a: 8 x 32-bit little-endian limbs in registers,
// products as IMAD.WIDE.U32(.X) carry chains (ptxas fuses each mad.lo.cc/madc.hi.cc pair into one
// IMAD.WIDE.U32.X with a predicate carry), adds as IADD3(.X) chains.
//
// Replaces the reference's WGSL field layer (src/shaders/field.wgsl:18-210: 16-bit split mul32,
// fold_single, fe_add/sub/mul/square/inv) — same mathematics, different machine mapping.
//
// Invariant: every fe is a 256-bit value in [0, 2^256) congruent to the field element ("weakly
// reduced"); fe_canon() brings it to [0, p). All routines are strictly correct for weak inputs.
//
// The same header compiles for the host with VG_HOST_EMU defined (carry flag emulated in a
// variable) so tests/ can exercise this exact code on a machine without a GPU.
#pragma once
#include <stdint.h>

#ifdef VG_HOST_EMU
#define VG_DEV static inline
#define VG_DEVM inline                 /* member functions */
#define VG_COLD static inline
static thread_local uint32_t vg_cc = 0;
static inline uint32_t vg_add_cc(uint32_t a, uint32_t b, uint32_t cin) { uint64_t t = (uint64_t)a + b + cin; vg_cc = (uint32_t)(t >> 32); return (uint32_t)t; }
static inline uint32_t vg_sub_cc(uint32_t a, uint32_t b, uint32_t bin) { uint64_t t = (uint64_t)a - b - bin; vg_cc = (uint32_t)(t >> 63); return (uint32_t)t; }
static inline uint32_t vg_mad_lo_cc(uint32_t a, uint32_t b, uint32_t c, uint32_t cin) { return vg_add_cc((uint32_t)((uint64_t)a * b), c, cin); }
static inline uint32_t vg_mad_hi_cc(uint32_t a, uint32_t b, uint32_t c, uint32_t cin) { return vg_add_cc((uint32_t)(((uint64_t)a * b) >> 32), c, cin); }
#define ADD_CC(r, a, b) { r = vg_add_cc(a, b, 0); }
#define ADDC_CC(r, a, b) { r = vg_add_cc(a, b, vg_cc); }
#define ADDC(r, a, b) { uint32_t vg_keep = vg_cc; r = vg_add_cc(a, b, vg_cc); vg_cc = vg_keep; }
#define SUB_CC(r, a, b) { r = vg_sub_cc(a, b, 0); }
#define SUBC_CC(r, a, b) { r = vg_sub_cc(a, b, vg_cc); }
#define SUBC(r, a, b) { uint32_t vg_keep = vg_cc; r = vg_sub_cc(a, b, vg_cc); vg_cc = vg_keep; }
#define MAD_LO_CC(r, a, b, c) { r = vg_mad_lo_cc(a, b, c, 0); }
#define MADC_LO_CC(r, a, b, c) { r = vg_mad_lo_cc(a, b, c, vg_cc); }
#define MADC_HI_CC(r, a, b, c) { r = vg_mad_hi_cc(a, b, c, vg_cc); }
#define MADC_HI(r, a, b, c) { uint32_t vg_keep = vg_cc; r = vg_mad_hi_cc(a, b, c, vg_cc); vg_cc = vg_keep; }
#define VG_UNROLL
#else
#define VG_DEV __device__ __forceinline__
#define VG_DEVM __device__ __forceinline__
#define VG_COLD static __device__ __noinline__
#define ADD_CC(r, a, b) asm volatile("add.cc.u32 %0,%1,%2;" : "=r"(r) : "r"(a), "r"(b))
#define ADDC_CC(r, a, b) asm volatile("addc.cc.u32 %0,%1,%2;" : "=r"(r) : "r"(a), "r"(b))
#define ADDC(r, a, b) asm volatile("addc.u32 %0,%1,%2;" : "=r"(r) : "r"(a), "r"(b))
#define SUB_CC(r, a, b) asm volatile("sub.cc.u32 %0,%1,%2;" : "=r"(r) : "r"(a), "r"(b))
#define SUBC_CC(r, a, b) asm volatile("subc.cc.u32 %0,%1,%2;" : "=r"(r) : "r"(a), "r"(b))
#define SUBC(r, a, b) asm volatile("subc.u32 %0,%1,%2;" : "=r"(r) : "r"(a), "r"(b))
#define MAD_LO_CC(r, a, b, c) asm volatile("mad.lo.cc.u32 %0,%1,%2,%3;" : "=r"(r) : "r"(a), "r"(b), "r"(c))
#define MADC_LO_CC(r, a, b, c) asm volatile("madc.lo.cc.u32 %0,%1,%2,%3;" : "=r"(r) : "r"(a), "r"(b), "r"(c))
#define MADC_HI_CC(r, a, b, c) asm volatile("madc.hi.cc.u32 %0,%1,%2,%3;" : "=r"(r) : "r"(a), "r"(b), "r"(c))
#define MADC_HI(r, a, b, c) asm volatile("madc.hi.u32 %0,%1,%2,%3;" : "=r"(r) : "r"(a), "r"(b), "r"(c))
#define VG_UNROLL _Pragma("unroll")
#endif

#define FE_RC_LO 977u   // 2^256 mod p = 2^32 + 977

struct fe { uint32_t v[8]; };

VG_DEV void fe_set_zero(fe &r) { VG_UNROLL for (int i = 0; i < 8; i++) r.v[i] = 0; }
VG_DEV void fe_set_one(fe &r) { fe_set_zero(r); r.v[0] = 1; }
VG_DEV bool fe_is_zero_raw(const fe &a) {   // raw 256-bit value == 0 (use on canonical values)
    return (a.v[0] | a.v[1] | a.v[2] | a.v[3] | a.v[4] | a.v[5] | a.v[6] | a.v[7]) == 0;
}
VG_DEV bool fe_eq_raw(const fe &a, const fe &b) {
    uint32_t d = 0;
    VG_UNROLL for (int i = 0; i < 8; i++) d |= a.v[i] ^ b.v[i];
    return d == 0;
}
// value >= p  <=>  limbs 7..2 all ones, limb1 == 0xFFFFFFFE.. : test cheaply, subtract rarely
VG_DEV void fe_canon(fe &a) {
    uint32_t t = a.v[7] & a.v[6] & a.v[5] & a.v[4] & a.v[3] & a.v[2];
    if (t == 0xFFFFFFFFu) {
        // p = 2^256 - 2^32 - 977: a >= p  <=>  a + 2^32 + 977 overflows 2^256
        uint32_t r[8], c;
        ADD_CC(r[0], a.v[0], FE_RC_LO); ADDC_CC(r[1], a.v[1], 1u);
        VG_UNROLL for (int i = 2; i < 8; i++) ADDC_CC(r[i], a.v[i], 0u);
        ADDC(c, 0u, 0u);
        if (c) { VG_UNROLL for (int i = 0; i < 8; i++) a.v[i] = r[i]; }
    }
}
// r = a + b
VG_DEV void fe_add(fe &r, const fe &a, const fe &b) {
    uint32_t c;
    ADD_CC(r.v[0], a.v[0], b.v[0]);
    VG_UNROLL for (int i = 1; i < 8; i++) ADDC_CC(r.v[i], a.v[i], b.v[i]);
    ADDC(c, 0u, 0u);
    // wrapped by 2^256 == RC (mod p): add c*RC
    uint32_t m = 0u - c;
    ADD_CC(r.v[0], r.v[0], m & FE_RC_LO); ADDC_CC(r.v[1], r.v[1], c);
    VG_UNROLL for (int i = 2; i < 8; i++) ADDC_CC(r.v[i], r.v[i], 0u);
    ADDC(c, 0u, 0u);
    if (c) {   // only when both inputs were within 2^33 of 2^256; result is tiny, no further carry
        ADD_CC(r.v[0], r.v[0], FE_RC_LO); ADDC_CC(r.v[1], r.v[1], 1u); ADDC(r.v[2], r.v[2], 0u);
    }
}
// r = a - b
VG_DEV void fe_sub(fe &r, const fe &a, const fe &b) {
    uint32_t m;
    SUB_CC(r.v[0], a.v[0], b.v[0]);
    VG_UNROLL for (int i = 1; i < 8; i++) SUBC_CC(r.v[i], a.v[i], b.v[i]);
    SUBC(m, 0u, 0u);                       // 0xFFFFFFFF on borrow
    // wrapped value = a - b + 2^256 == a - b + RC: subtract RC
    SUB_CC(r.v[0], r.v[0], m & FE_RC_LO); SUBC_CC(r.v[1], r.v[1], m & 1u);
    VG_UNROLL for (int i = 2; i < 8; i++) SUBC_CC(r.v[i], r.v[i], 0u);
    SUBC(m, 0u, 0u);
    if (m) {   // only when b was within 2^33 of 2^256
        SUB_CC(r.v[0], r.v[0], FE_RC_LO); SUBC_CC(r.v[1], r.v[1], 1u);
        VG_UNROLL for (int i = 2; i < 8; i++) SUBC_CC(r.v[i], r.v[i], 0u);
    }
}
VG_DEV void fe_neg(fe &r, const fe &a) { fe z; fe_set_zero(z); fe_sub(r, z, a); }

// acc[0..7] += {a[0], a[2], a[4], a[6]} * bi at 64-bit aligned pairs; leaves the carry in CC
#define VG_CHAIN4(acc, a, bi)                                                        \
    MAD_LO_CC((acc)[0], (a)[0], bi, (acc)[0]); MADC_HI_CC((acc)[1], (a)[0], bi, (acc)[1]); \
    MADC_LO_CC((acc)[2], (a)[2], bi, (acc)[2]); MADC_HI_CC((acc)[3], (a)[2], bi, (acc)[3]); \
    MADC_LO_CC((acc)[4], (a)[4], bi, (acc)[4]); MADC_HI_CC((acc)[5], (a)[4], bi, (acc)[5]); \
    MADC_LO_CC((acc)[6], (a)[6], bi, (acc)[6]); MADC_HI_CC((acc)[7], (a)[6], bi, (acc)[7]);

// 256 x 256 -> 512: products with even (i+j) accumulate in ev[], odd in od[] (od[k] sits at limb k+1),
// so every product lands on an aligned register pair = one IMAD.WIDE.U32.X each (64 total).
VG_DEV void fe_mul_wide(uint32_t r[16], const uint32_t a[8], const uint32_t b[8]) {
    uint32_t ev[16], od[16];
    VG_UNROLL for (int i = 0; i < 16; i++) { ev[i] = 0; od[i] = 0; }
    VG_UNROLL for (int i = 0; i < 8; i += 2) {
        VG_CHAIN4(ev + i, a, b[i]);          if (i) ADDC(ev[i + 8], ev[i + 8], 0u);
        VG_CHAIN4(od + i, a + 1, b[i]);      if (i) ADDC(od[i + 8], od[i + 8], 0u);
        VG_CHAIN4(od + i, a, b[i + 1]);      ADDC(od[i + 8], od[i + 8], 0u);
        VG_CHAIN4(ev + i + 2, a + 1, b[i + 1]);
    }
    r[0] = ev[0];
    ADD_CC(r[1], ev[1], od[0]);
    VG_UNROLL for (int k = 2; k < 15; k++) ADDC_CC(r[k], ev[k], od[k - 1]);
    ADDC(r[15], ev[15], od[14]);
}
// r = t mod p (weakly): fold the high 256 bits by 2^256 == 2^32 + 977, twice, then once more.
VG_DEV void fe_reduce_wide(fe &r, const uint32_t t[16]) {
    uint32_t lo[10];
    VG_UNROLL for (int i = 0; i < 8; i++) lo[i] = t[i];
    lo[8] = 0; lo[9] = 0;
    const uint32_t *h = t + 8;
    // + hi * 977 (even / odd aligned chains)
    MAD_LO_CC(lo[0], h[0], FE_RC_LO, lo[0]); MADC_HI_CC(lo[1], h[0], FE_RC_LO, lo[1]);
    MADC_LO_CC(lo[2], h[2], FE_RC_LO, lo[2]); MADC_HI_CC(lo[3], h[2], FE_RC_LO, lo[3]);
    MADC_LO_CC(lo[4], h[4], FE_RC_LO, lo[4]); MADC_HI_CC(lo[5], h[4], FE_RC_LO, lo[5]);
    MADC_LO_CC(lo[6], h[6], FE_RC_LO, lo[6]); MADC_HI_CC(lo[7], h[6], FE_RC_LO, lo[7]);
    ADDC(lo[8], lo[8], 0u);
#if defined(VG_REDUCE_SPLIT) && VG_REDUCE_SPLIT
    {   // the odd limbs' products in their own aligned pairs (no addend, no carry: four independent wide products), then ONE add chain
        // shifted by a limb. Accumulating them straight into lo[1..8] makes ptxas copy lo[] into pairs of the other alignment and
        // back (16 IMAD.MOV on the multiplier pipe per reduction). Measured on the P2TR stage, where that pipe binds: 1,045 fewer
        // IMAD.MOV in the kernel, yet 872 against 878 Mkeys/s (profiles/reduce_split_r02.log), so the switch stays off.
        uint32_t od[8];
        VG_UNROLL for (int k = 0; k < 4; k++) { MAD_LO_CC(od[2 * k], h[2 * k + 1], FE_RC_LO, 0u); MADC_HI(od[2 * k + 1], h[2 * k + 1], FE_RC_LO, 0u); }
        ADD_CC(lo[1], lo[1], od[0]);
        VG_UNROLL for (int k = 1; k < 8; k++) ADDC_CC(lo[k + 1], lo[k + 1], od[k]);
        ADDC(lo[9], lo[9], 0u);
    }
#else
    MAD_LO_CC(lo[1], h[1], FE_RC_LO, lo[1]); MADC_HI_CC(lo[2], h[1], FE_RC_LO, lo[2]);
    MADC_LO_CC(lo[3], h[3], FE_RC_LO, lo[3]); MADC_HI_CC(lo[4], h[3], FE_RC_LO, lo[4]);
    MADC_LO_CC(lo[5], h[5], FE_RC_LO, lo[5]); MADC_HI_CC(lo[6], h[5], FE_RC_LO, lo[6]);
    MADC_LO_CC(lo[7], h[7], FE_RC_LO, lo[7]); MADC_HI_CC(lo[8], h[7], FE_RC_LO, lo[8]);
    ADDC(lo[9], lo[9], 0u);
#endif
    // + hi * 2^32
    ADD_CC(lo[1], lo[1], h[0]);
    VG_UNROLL for (int i = 1; i < 8; i++) ADDC_CC(lo[i + 1], lo[i + 1], h[i]);
    ADDC(lo[9], lo[9], 0u);
    // second fold: C = lo[8] + 2^32*lo[9] < 2^34;  F = C*977 + C*2^32 < 2^67
    uint32_t f0, f1, f2, f3, c;
    MAD_LO_CC(f0, lo[8], FE_RC_LO, 0u); MADC_HI_CC(f1, lo[8], FE_RC_LO, 0u); ADDC(f2, 0u, 0u);
    MAD_LO_CC(f1, lo[9], FE_RC_LO, f1); MADC_HI_CC(f2, lo[9], FE_RC_LO, f2); ADDC(f3, 0u, 0u);
    ADD_CC(f1, f1, lo[8]); ADDC_CC(f2, f2, lo[9]); ADDC(f3, f3, 0u);
    ADD_CC(r.v[0], lo[0], f0); ADDC_CC(r.v[1], lo[1], f1); ADDC_CC(r.v[2], lo[2], f2); ADDC_CC(r.v[3], lo[3], f3);
    VG_UNROLL for (int i = 4; i < 8; i++) ADDC_CC(r.v[i], lo[i], 0u);
    ADDC(c, 0u, 0u);
    // third fold: only if the sum wrapped, and then the value is < 2^68, so no carry past limb 3
    uint32_t m = 0u - c;
    ADD_CC(r.v[0], r.v[0], m & FE_RC_LO); ADDC_CC(r.v[1], r.v[1], c); ADDC_CC(r.v[2], r.v[2], 0u); ADDC(r.v[3], r.v[3], 0u);
}
VG_DEV void fe_mul(fe &r, const fe &a, const fe &b) {
    uint32_t t[16];
    fe_mul_wide(t, a.v, b.v);
    fe_reduce_wide(r, t);
}
// 256-bit square -> 512 bits with 36 wide products instead of 64: the 28 cross products a_i*a_j (i < j) once, in the same
// even/odd aligned accumulators as fe_mul_wide (rows taken in order of j so that every chain's carry lands in a limb
// nothing has written yet), doubled, plus the 8 squares on one carry chain. Costs ~25 more add-class instructions
// than fe_mul_wide, so it pays where the multiplier pipe binds (field inversion, the P2TR tweak stage) and not in the walk
// kernel's per-key lambda^2, where the ALU pipe binds (fe_sqr stays a plain multiplication there).
VG_DEV void fe_sqr_wide(uint32_t r[16], const uint32_t a[8]) {
    uint32_t ev[16], od[16];
    VG_UNROLL for (int i = 0; i < 16; i++) { ev[i] = 0; od[i] = 0; }
    MAD_LO_CC(od[0], a[0], a[1], od[0]); MADC_HI_CC(od[1], a[0], a[1], od[1]); ADDC(od[2], od[2], 0u);
    MAD_LO_CC(ev[2], a[0], a[2], ev[2]); MADC_HI_CC(ev[3], a[0], a[2], ev[3]); ADDC(ev[4], ev[4], 0u);
    MAD_LO_CC(od[2], a[1], a[2], od[2]); MADC_HI_CC(od[3], a[1], a[2], od[3]); ADDC(od[4], od[4], 0u);
    MAD_LO_CC(ev[4], a[1], a[3], ev[4]); MADC_HI_CC(ev[5], a[1], a[3], ev[5]); ADDC(ev[6], ev[6], 0u);
    MAD_LO_CC(od[2], a[0], a[3], od[2]); MADC_HI_CC(od[3], a[0], a[3], od[3]); MADC_LO_CC(od[4], a[2], a[3], od[4]); MADC_HI_CC(od[5], a[2], a[3], od[5]); ADDC(od[6], od[6], 0u);
    MAD_LO_CC(ev[4], a[0], a[4], ev[4]); MADC_HI_CC(ev[5], a[0], a[4], ev[5]); MADC_LO_CC(ev[6], a[2], a[4], ev[6]); MADC_HI_CC(ev[7], a[2], a[4], ev[7]); ADDC(ev[8], ev[8], 0u);
    MAD_LO_CC(od[4], a[1], a[4], od[4]); MADC_HI_CC(od[5], a[1], a[4], od[5]); MADC_LO_CC(od[6], a[3], a[4], od[6]); MADC_HI_CC(od[7], a[3], a[4], od[7]); ADDC(od[8], od[8], 0u);
    MAD_LO_CC(ev[6], a[1], a[5], ev[6]); MADC_HI_CC(ev[7], a[1], a[5], ev[7]); MADC_LO_CC(ev[8], a[3], a[5], ev[8]); MADC_HI_CC(ev[9], a[3], a[5], ev[9]); ADDC(ev[10], ev[10], 0u);
    MAD_LO_CC(od[4], a[0], a[5], od[4]); MADC_HI_CC(od[5], a[0], a[5], od[5]); MADC_LO_CC(od[6], a[2], a[5], od[6]); MADC_HI_CC(od[7], a[2], a[5], od[7]); MADC_LO_CC(od[8], a[4], a[5], od[8]); MADC_HI_CC(od[9], a[4], a[5], od[9]); ADDC(od[10], od[10], 0u);
    MAD_LO_CC(ev[6], a[0], a[6], ev[6]); MADC_HI_CC(ev[7], a[0], a[6], ev[7]); MADC_LO_CC(ev[8], a[2], a[6], ev[8]); MADC_HI_CC(ev[9], a[2], a[6], ev[9]); MADC_LO_CC(ev[10], a[4], a[6], ev[10]); MADC_HI_CC(ev[11], a[4], a[6], ev[11]); ADDC(ev[12], ev[12], 0u);
    MAD_LO_CC(od[6], a[1], a[6], od[6]); MADC_HI_CC(od[7], a[1], a[6], od[7]); MADC_LO_CC(od[8], a[3], a[6], od[8]); MADC_HI_CC(od[9], a[3], a[6], od[9]); MADC_LO_CC(od[10], a[5], a[6], od[10]); MADC_HI_CC(od[11], a[5], a[6], od[11]); ADDC(od[12], od[12], 0u);
    MAD_LO_CC(ev[8], a[1], a[7], ev[8]); MADC_HI_CC(ev[9], a[1], a[7], ev[9]); MADC_LO_CC(ev[10], a[3], a[7], ev[10]); MADC_HI_CC(ev[11], a[3], a[7], ev[11]); MADC_LO_CC(ev[12], a[5], a[7], ev[12]); MADC_HI_CC(ev[13], a[5], a[7], ev[13]); ADDC(ev[14], ev[14], 0u);
    MAD_LO_CC(od[6], a[0], a[7], od[6]); MADC_HI_CC(od[7], a[0], a[7], od[7]); MADC_LO_CC(od[8], a[2], a[7], od[8]); MADC_HI_CC(od[9], a[2], a[7], od[9]); MADC_LO_CC(od[10], a[4], a[7], od[10]); MADC_HI_CC(od[11], a[4], a[7], od[11]); MADC_LO_CC(od[12], a[6], a[7], od[12]); MADC_HI_CC(od[13], a[6], a[7], od[13]); ADDC(od[14], od[14], 0u);
    // cross = ev + (od << 32)
    r[0] = ev[0];
    ADD_CC(r[1], ev[1], od[0]);
    VG_UNROLL for (int k = 2; k < 15; k++) ADDC_CC(r[k], ev[k], od[k - 1]);
    ADDC(r[15], ev[15], od[14]);
    // 2 * cross (< 2^512)
    ADD_CC(r[0], r[0], r[0]);
    VG_UNROLL for (int k = 1; k < 15; k++) ADDC_CC(r[k], r[k], r[k]);
    ADDC(r[15], r[15], r[15]);
    // + a_i^2 at limb 2i
    MAD_LO_CC(r[0], a[0], a[0], r[0]); MADC_HI_CC(r[1], a[0], a[0], r[1]);
    VG_UNROLL for (int i = 1; i < 7; i++) { MADC_LO_CC(r[2 * i], a[i], a[i], r[2 * i]); MADC_HI_CC(r[2 * i + 1], a[i], a[i], r[2 * i + 1]); }
    MADC_LO_CC(r[14], a[7], a[7], r[14]); MADC_HI(r[15], a[7], a[7], r[15]);
}
VG_DEV void fe_sqr_fast(fe &r, const fe &a) {
    uint32_t t[16];
    fe_sqr_wide(t, a.v);
    fe_reduce_wide(r, t);
}
VG_DEV void fe_sqr(fe &r, const fe &a) { fe_mul(r, a, a); }
VG_DEV void fe_sqr_n(fe &r, const fe &a, int n) {
    r = a;
#if defined(VG_INV_FAST_SQR) && VG_INV_FAST_SQR
    for (int i = 0; i < n; i++) fe_sqr_fast(r, r);     // P2TR build: the multiplier pipe binds (one inversion per 16 keys)
#else
    for (int i = 0; i < n; i++) fe_sqr(r, r);          // walk kernels: the ALU pipe binds; measured 6,288 vs 6,268 Mkeys/s
#endif
}
// a^(p-2); p-2 = (223 ones) 0 (22 ones) 0000 1 0 11 0 1 in binary. 255 squarings + 15 multiplies.
VG_COLD void fe_inv(fe &r, const fe &a) {
    fe x2, x3, x6, x9, x11, x22, x44, x88, x176, x220, x223, t;
    fe_sqr(t, a); fe_mul(x2, t, a);
    fe_sqr(t, x2); fe_mul(x3, t, a);
    fe_sqr_n(t, x3, 3); fe_mul(x6, t, x3);
    fe_sqr_n(t, x6, 3); fe_mul(x9, t, x3);
    fe_sqr_n(t, x9, 2); fe_mul(x11, t, x2);
    fe_sqr_n(t, x11, 11); fe_mul(x22, t, x11);
    fe_sqr_n(t, x22, 22); fe_mul(x44, t, x22);
    fe_sqr_n(t, x44, 44); fe_mul(x88, t, x44);
    fe_sqr_n(t, x88, 88); fe_mul(x176, t, x88);
    fe_sqr_n(t, x176, 44); fe_mul(x220, t, x44);
    fe_sqr_n(t, x220, 3); fe_mul(x223, t, x3);
    fe_sqr_n(t, x223, 23); fe_mul(t, t, x22);
    fe_sqr_n(t, t, 5); fe_mul(t, t, a);
    fe_sqr_n(t, t, 3); fe_mul(t, t, x2);
    fe_sqr_n(t, t, 2); fe_mul(r, t, a);
}
