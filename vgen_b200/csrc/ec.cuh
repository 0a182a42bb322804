// secp256k1 group operations for the scan kernels.
//
// Hot path (walk kernel): affine addition with a supplied inverse — 2M + 1S per point — fed by a
// Montgomery batch inversion over each thread's key batch (scan_kernels.cu). This replaces the
// reference's per-key Jacobian add + Fermat inversion (src/shaders/search.wgsl:2-31,
// src/shaders/field.wgsl:195-210,234-342).
// Cold path (table / centre set-up, degenerate batches): Jacobian double / mixed add and a
// table-driven scalar multiplication (replaces scalar_mul_G, src/shaders/field.wgsl:352-363 and the
// host-side key_to_affine, src/gpu.rs:901-910).
#pragma once
#include "fe.cuh"

struct gej_t { fe x, y, z; bool inf; };

// secp256k1 group order n, little-endian 32-bit limbs
#define VG_N0 0xD0364141u
#define VG_N1 0xBFD25E8Cu
#define VG_N2 0xAF48A03Bu
#define VG_N3 0xBAAEDCE6u
#define VG_N4 0xFFFFFFFEu
#define VG_N5 0xFFFFFFFFu
#define VG_N6 0xFFFFFFFFu
#define VG_N7 0xFFFFFFFFu

// 0 < k < n ?
VG_DEV bool scalar_is_valid(const uint32_t k[8]) {
    const uint32_t n[8] = {VG_N0, VG_N1, VG_N2, VG_N3, VG_N4, VG_N5, VG_N6, VG_N7};
    uint32_t any = 0;
    VG_UNROLL for (int i = 0; i < 8; i++) any |= k[i];
    if (!any) return false;
    for (int i = 7; i >= 0; i--) {
        if (k[i] < n[i]) return true;
        if (k[i] > n[i]) return false;
    }
    return false;
}
// r = a + v (v < 2^64), wrapping at 2^256
VG_DEV void scalar_add_u64(uint32_t r[8], const uint32_t a[8], uint64_t v) {
    ADD_CC(r[0], a[0], (uint32_t)v); ADDC_CC(r[1], a[1], (uint32_t)(v >> 32));
    VG_UNROLL for (int i = 2; i < 8; i++) ADDC_CC(r[i], a[i], 0u);
}

VG_COLD void gej_double(gej_t &r, const gej_t &a) {
    fe ycan = a.y; fe_canon(ycan);
    if (a.inf || fe_is_zero_raw(ycan)) { r.inf = true; return; }
    fe A, B, C, D, E, F, t, x3, y3, z3;
    fe_sqr(A, a.x); fe_sqr(B, a.y); fe_sqr(C, B);
    fe_add(t, a.x, B); fe_sqr(t, t); fe_sub(t, t, A); fe_sub(t, t, C);
    fe_add(D, t, t);
    fe_add(E, A, A); fe_add(E, E, A);
    fe_sqr(F, E);
    fe_sub(x3, F, D); fe_sub(x3, x3, D);
    fe_sub(t, D, x3); fe_mul(y3, E, t);
    fe_add(t, C, C); fe_add(t, t, t); fe_add(t, t, t);
    fe_sub(y3, y3, t);
    fe_mul(z3, a.y, a.z); fe_add(z3, z3, z3);
    r.x = x3; r.y = y3; r.z = z3; r.inf = false;
}
// r = a + (bx, by), all degenerate cases handled
VG_COLD void gej_add_affine(gej_t &r, const gej_t &a, const fe &bx, const fe &by) {
    if (a.inf) { r.x = bx; r.y = by; fe_set_one(r.z); r.inf = false; return; }
    fe z2, u2, s2, h, rr, h2, h3, v, t, x3, y3, z3;
    fe_sqr(z2, a.z); fe_mul(u2, bx, z2);
    fe_mul(s2, by, z2); fe_mul(s2, s2, a.z);
    fe_sub(h, u2, a.x); fe_sub(rr, s2, a.y);
    fe_canon(h);
    if (fe_is_zero_raw(h)) {
        fe_canon(rr);
        if (fe_is_zero_raw(rr)) { gej_t d = a; gej_double(r, d); return; }
        r.inf = true; return;
    }
    fe_sqr(h2, h); fe_mul(h3, h2, h); fe_mul(v, a.x, h2);
    fe_sqr(x3, rr); fe_sub(x3, x3, h3); fe_sub(x3, x3, v); fe_sub(x3, x3, v);
    fe_sub(t, v, x3); fe_mul(y3, rr, t); fe_mul(t, a.y, h3); fe_sub(y3, y3, t);
    fe_mul(z3, a.z, h);
    r.x = x3; r.y = y3; r.z = z3; r.inf = false;
}
// canonical affine coordinates; false if the point is at infinity
VG_COLD bool gej_to_affine(fe &x, fe &y, const gej_t &a) {
    if (a.inf) { fe_set_zero(x); fe_set_zero(y); return false; }
    fe zi, zi2, zi3;
    fe_inv(zi, a.z); fe_sqr(zi2, zi); fe_mul(zi3, zi2, zi);
    fe_mul(x, a.x, zi2); fe_mul(y, a.y, zi3);
    fe_canon(x); fe_canon(y);
    return true;
}
// k*G from the table pow2[i] = 2^i * G (affine, canonical; x then y, 8 limbs each, 256 entries).
// Returns false when the result is the point at infinity (k == 0 mod n).
VG_COLD bool scalar_mul_pow2(fe &x, fe &y, const uint32_t k[8], const uint32_t *pow2) {
    gej_t acc; acc.inf = true;
    fe_set_zero(acc.x); fe_set_zero(acc.y); fe_set_zero(acc.z);
    for (int i = 0; i < 256; i++) {
        if ((k[i >> 5] >> (i & 31)) & 1u) {
            fe bx, by;
            VG_UNROLL for (int j = 0; j < 8; j++) { bx.v[j] = pow2[i * 16 + j]; by.v[j] = pow2[i * 16 + 8 + j]; }
            gej_t nr; gej_add_affine(nr, acc, bx, by); acc = nr;
        }
    }
    return gej_to_affine(x, y, acc);
}
// Generator tables in one device buffer: pow2[256][16] (2^i * G) followed by win8[32][256][16] (v * 2^(8j) * G, v = 0 unused).
#define VG_GTABLE_POW2_WORDS (256 * 16)
#ifndef VG_GTABLE_WORDS
#define VG_GTABLE_WORDS (VG_GTABLE_POW2_WORDS + 32 * 256 * 16)
#endif
// k*G by 8-bit windows: at most 32 additions (the bit-by-bit table above needs ~128). Used for the centre set-up of every
// scan (one scalar multiplication per thread, replacing the host-side key_to_affine of src/gpu.rs:901-910) and for the
// degenerate-batch path. w8 = false falls back to the pow2 table (host emulation builds only that part by default).
VG_COLD bool scalar_mul_G(fe &x, fe &y, const uint32_t k[8], const uint32_t *gtable, bool w8) {
    if (!w8) return scalar_mul_pow2(x, y, k, gtable);
    const uint32_t *win8 = gtable + VG_GTABLE_POW2_WORDS;
    gej_t acc; acc.inf = true;
    fe_set_zero(acc.x); fe_set_zero(acc.y); fe_set_zero(acc.z);
    for (int j = 0; j < 32; j++) {
        const uint32_t v = (k[j >> 2] >> (8 * (j & 3))) & 0xffu;
        if (!v) continue;
        const uint32_t *e = win8 + ((size_t)j * 256 + v) * 16;
        fe bx, by;
        VG_UNROLL for (int l = 0; l < 8; l++) { bx.v[l] = e[l]; by.v[l] = e[8 + l]; }
        gej_t nr; gej_add_affine(nr, acc, bx, by); acc = nr;
    }
    return gej_to_affine(x, y, acc);
}
#ifdef VG_HOST_EMU
#define VG_W8 false
#else
#define VG_W8 true
#endif
// affine add given inv = 1/(bx - ax): (x3, y3) = (ax, ay) + (bx, by). 2M + 1S. Outputs weakly reduced.
VG_DEV void affine_add_inv(fe &x3, fe &y3, const fe &ax, const fe &ay, const fe &bx, const fe &by, const fe &inv) {
    fe dy, lam, t;
    fe_sub(dy, by, ay);
    fe_mul(lam, dy, inv);
    fe_sqr(x3, lam); fe_sub(x3, x3, ax); fe_sub(x3, x3, bx);
    fe_sub(t, ax, x3); fe_mul(y3, lam, t); fe_sub(y3, y3, ay);
}
