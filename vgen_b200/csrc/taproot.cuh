// P2TR (BIP341 key-path output, no script tree) on the device: second stage of the scan for VG_FMT_P2TR.
//
// The reference's GPU path stops at the even-Y internal key (src/shaders/search_p2tr.wgsl:104-113) and leaves the
// TapTweak to the host: per key a tagged SHA-256 and a full scalar multiplication through
// bitcoin::Address::p2tr (src/gpu.rs:1288-1290, src/address.rs:136-140). Here the whole derivation runs on the GPU:
//
//   stage 1 (walk kernel, FMT = P2TR): writes the public key (x, y) of every key of the window into
//           `points` (64 B per key, limb order as `fe`), all-zero for keys that are invalid / outside the range;
//   stage 2 (this file): t = SHA256(SHA256("TapTweak") || SHA256("TapTweak") || x)   [one compression from a
//           precomputed midstate, same constant as src/shaders/sha256.wgsl:177-249]
//           Q = P_even + t*G. The key's scalar k is known (window base + index), and P_even = k*G for even y, (n - k)*G for odd, so
//           Q = s*G with s = t + k or t + n - k (mod n): ONE fixed-base multiplication, no addition of P (VG_TAP_SCALAR; with it off,
//           stage 1 stores the even-Y point and stage 2 adds t*G to it: one table addition more per key).
//           The multiplication runs through a fixed-base table of affine multiples v * 2^(off_j) * G per window j (geometry word:
//           scan_types.h; signed-digit windows: 15 of them in 67 MB for short jobs, 10 in 17 GB, 9 in 111.7 GB of HBM for the longest,
//           gathered at random), accumulated in XYZZ coordinates; one field inversion per VG_TAP_GROUP keys (Montgomery's
//           trick over the group's ZZ = Z^2) turns X/ZZ into the 32-byte witness program x(Q), which goes through the
//           same matcher as the other formats (interval / exact / Bech32m-fed DFA).
#pragma once
#include "match.cuh"

#ifndef VG_TAP_SYNC
#define VG_TAP_SYNC 1
#endif
#ifndef VG_TAP_GROUP
#define VG_TAP_GROUP 64         // keys handled by one stage-2 thread = keys sharing one field inversion (<= 64; measured 16: 684, 32: 722, 64: 742 Mkeys/s)
#endif

// t as 8 big-endian words (t[0] most significant) of the TapTweak tagged hash of the x-only key
VG_DEV void vg_taptweak_hash(uint32_t t[8], const fe &x) {
    const uint32_t mid[8] = {0xd129a2f3u, 0x701c655du, 0x6583b6c3u, 0xb9419727u, 0x95f4e232u, 0x94fd54f4u, 0xa2ae8d85u, 0x47ca590bu};
    uint32_t w[16];
    VG_UNROLL for (int i = 0; i < 8; i++) { t[i] = mid[i]; w[i] = x.v[7 - i]; }
    w[8] = 0x80000000u;
    VG_UNROLL for (int i = 9; i < 15; i++) w[i] = 0;
    w[15] = 96 * 8;
    sha256_compress<0xFF00u>(t, w);
}
// t < n ?  (t big-endian words)
VG_DEV bool vg_tweak_is_valid(const uint32_t t[8]) {
    uint32_t k[8];
    VG_UNROLL for (int i = 0; i < 8; i++) k[i] = t[7 - i];
    uint32_t any = 0;
    VG_UNROLL for (int i = 0; i < 8; i++) any |= k[i];
    return any == 0 || scalar_is_valid(k);          // t == 0 is allowed by BIP341 (Q = P)
}
// r = n - k for 0 < k < n (little-endian limbs)
VG_DEV void vg_scalar_neg_n(uint32_t r[8], const uint32_t k[8]) {
    const uint32_t n[8] = {VG_N0, VG_N1, VG_N2, VG_N3, VG_N4, VG_N5, VG_N6, VG_N7};
    SUB_CC(r[0], n[0], k[0]);
    VG_UNROLL for (int i = 1; i < 7; i++) SUBC_CC(r[i], n[i], k[i]);
    SUBC(r[7], n[7], k[7]);
}
// r = a + b mod n for a, b < n (little-endian limbs)
VG_DEV void vg_scalar_add_mod_n(uint32_t r[8], const uint32_t a[8], const uint32_t b[8]) {
    const uint32_t n[8] = {VG_N0, VG_N1, VG_N2, VG_N3, VG_N4, VG_N5, VG_N6, VG_N7};
    uint32_t s[8], d[8], c, br;
    ADD_CC(s[0], a[0], b[0]);
    VG_UNROLL for (int i = 1; i < 8; i++) ADDC_CC(s[i], a[i], b[i]);
    ADDC(c, 0u, 0u);                                   // a + b >= 2^256
    SUB_CC(d[0], s[0], n[0]);
    VG_UNROLL for (int i = 1; i < 8; i++) SUBC_CC(d[i], s[i], n[i]);
    SUBC(br, 0u, 0u);                                  // 0xFFFFFFFF when s < n
    const bool take = c != 0 || br == 0;               // the true sum is >= n
    VG_UNROLL for (int i = 0; i < 8; i++) r[i] = take ? d[i] : s[i];
}
// r = a + (bx, by) for a finite Jacobian a, generic case only (8M + 3S, no calls). Returns false when the two points share
// their x coordinate (doubling / opposite points): the caller then redoes the key with the complete formulas.
// need_y = false (the last table window: only X and Z of the sum are used) skips the two multiplications of Y3.
VG_DEV bool vg_gej_add_affine_fast(gej_t &a, const fe &bx, const fe &by, bool need_y = true) {
    fe z2, u2, s2, h, rr, h2, h3, v, t;
    fe_sqr_fast(z2, a.z); fe_mul(u2, bx, z2);
    fe_mul(s2, by, z2); fe_mul(s2, s2, a.z);
    fe_sub(h, u2, a.x); fe_sub(rr, s2, a.y);
    fe hc = h; fe_canon(hc);
    if (fe_is_zero_raw(hc)) return false;
    fe_sqr_fast(h2, h); fe_mul(h3, h2, h); fe_mul(v, a.x, h2);
    fe_sqr_fast(t, rr); fe_sub(t, t, h3); fe_sub(t, t, v); fe_sub(t, t, v);      // X3
    if (need_y) { fe_sub(v, v, t); fe_mul(v, rr, v); fe_mul(h3, a.y, h3); fe_sub(a.y, v, h3); }
    a.x = t;
    fe_mul(a.z, a.z, h);
    return true;
}
// The same for a.z == 1 (the first table addition of a key: P is affine): 4M + 2S. VG_TAP_FIRST_AFFINE = 0 switches it off.
#ifndef VG_TAP_FIRST_AFFINE
#define VG_TAP_FIRST_AFFINE 1
#endif
// VG_TAP_LAST_NO_Y: the addition of the top table window leaves Y3 out (the caller reads only X and Z of Q)
#ifndef VG_TAP_LAST_NO_Y
#define VG_TAP_LAST_NO_Y 1
#endif
VG_DEV bool vg_aff_add_affine_fast(gej_t &a, const fe &bx, const fe &by) {
    fe h, rr, h2, h3, v, t;
    fe_sub(h, bx, a.x); fe_sub(rr, by, a.y);
    fe hc = h; fe_canon(hc);
    if (fe_is_zero_raw(hc)) return false;
    fe_sqr_fast(h2, h); fe_mul(h3, h2, h); fe_mul(v, a.x, h2);
    fe_sqr_fast(t, rr); fe_sub(t, t, h3); fe_sub(t, t, v); fe_sub(t, t, v);      // X3
    fe_sub(v, v, t); fe_mul(v, rr, v); fe_mul(h3, a.y, h3); fe_sub(a.y, v, h3);
    a.x = t;
    a.z = h;
    return true;
}

// Fixed-base table geometry: scan_types.h (vg_tap_windows / vg_tap_window / vg_tap_entries over the geometry word).
// k = v * 2^off as 8 little-endian limbs; false when v is 0 or the product does not fit in 256 bits (unused entry)
VG_DEV bool vg_tap_digit_scalar(uint32_t k[8], uint32_t off, uint32_t v) {
    const uint32_t wi = off >> 5, sh = off & 31u;
    for (int i = 0; i < 8; i++) k[i] = 0;
    if (v == 0 || off >= 256u) return false;
    if (off + 32u > 256u && (v >> (256u - off)) != 0) return false;
    k[wi] = v << sh;
    if (sh != 0 && wi + 1 < 8) k[wi + 1] = v >> (32u - sh);
    return true;
}
// the `width` bits of t from bit `off` up (t as 8 big-endian words: word 7 is the least significant; bits past 255 read as 0)
VG_DEV uint32_t vg_tap_digit(const uint32_t t[8], uint32_t off, uint32_t width) {
    const uint32_t wi = off >> 5, sh = off & 31u, mask = (width >= 32u) ? 0xffffffffu : ((1u << width) - 1u);
    uint32_t lo = 0, hi = 0;
    VG_UNROLL for (int i = 0; i < 8; i++) { lo = (wi == (uint32_t)i) ? t[7 - i] : lo; hi = (wi + 1u == (uint32_t)i) ? t[7 - i] : hi; }
    return (sh ? ((lo >> sh) | (hi << (32u - sh))) : lo) & mask;
}
// signed geometry: x(s*G) = x((n - s)*G) and only x(Q) leaves the stage, so the smaller of s and n - s is multiplied: it is below
// 2^255, which keeps the top window's signed digit inside its table (no carry out of bit 255). t: 8 big-endian words, 0 < t < n.
VG_DEV void vg_tap_fold_scalar(uint32_t t[8]) {
    uint32_t s[8], m[8];
    VG_UNROLL for (int i = 0; i < 8; i++) s[i] = t[7 - i];
    vg_scalar_neg_n(m, s);
    bool m_less = false, decided = false;
    VG_UNROLL for (int i = 7; i >= 0; i--) { if (!decided && m[i] != s[i]) { m_less = m[i] < s[i]; decided = true; } }
    VG_UNROLL for (int i = 0; i < 8; i++) t[7 - i] = m_less ? m[i] : s[i];
}

// Table entry of window j for the scalar tt (8 big-endian words): its index in the table, 0 when the digit is 0 (index 0 is the unused
// zero multiple of window 0). Signed geometries: `carry` runs from window to window, `neg` says the entry's y is to be negated.
VG_DEV unsigned long long vg_tap_entry(uint32_t g, const uint32_t tt[8], uint32_t j, uint32_t &carry, bool &neg) {
    uint32_t off, width; unsigned long long first;
    vg_tap_window(g, j, &off, &width, &first);
    uint32_t v = vg_tap_digit(tt, off, width) + carry;
    neg = false;
    if (g & VG_TAP_SIGNED) {                    // digit in [-2^(width-1), 2^(width-1)]: a value above the half borrows from the next window
        const uint32_t half = 1u << (width - 1u);
        carry = v > half ? 1u : 0u;
        if (carry) { v = (1u << width) - v; neg = true; }
    }
    return v ? first + v : 0ull;
}
// VG_TAP_PREFETCH = 1: before a key's additions start, all of its table entries are requested into L2 (they are gathered at random
// from a table of 43 to 112 GB; the additions of the earlier windows would then cover the DRAM latency of the later ones).
// Measured on B200 (profiles/tap_prefetch_r02.log): 719 against 901 Mkeys/s with the 9-window table, 791 against 856 with the 26-bit
// one — the stage is not waiting for those loads (16 warps per SM cover them) and the prefetches only add traffic: off.
#ifndef VG_TAP_PREFETCH
#define VG_TAP_PREFETCH 0
#endif

// The accumulator of the fast path in XYZZ coordinates (x = X / ZZ, y = Y / ZZZ, ZZ^3 = ZZZ^2): a mixed addition costs 8M + 2S
// instead of the Jacobian 8M + 3S, the top window (no Y3, no ZZZ3) 5M + 2S instead of 6M + 3S, and the caller gets ZZ = Z^2, which
// is what the grouped inversion wants, without a squaring. VG_TAP_XYZZ = 0: Jacobian accumulator.
#ifndef VG_TAP_XYZZ
#define VG_TAP_XYZZ 1
#endif
// (X, Y, ZZ, ZZZ) += (bx, by), generic case only; false when the two points share their x coordinate
VG_DEV bool vg_xyzz_add_affine_fast(fe &X, fe &Y, fe &ZZ, fe &ZZZ, const fe &bx, const fe &by, bool need_y) {
    fe u2, s2, p, r, pp, ppp, q, t;
    fe_mul(u2, bx, ZZ); fe_mul(s2, by, ZZZ);
    fe_sub(p, u2, X); fe_sub(r, s2, Y);
    fe pc = p; fe_canon(pc);
    if (fe_is_zero_raw(pc)) return false;
    fe_sqr_fast(pp, p); fe_mul(ppp, pp, p); fe_mul(q, X, pp);
    fe_sqr_fast(t, r); fe_sub(t, t, ppp); fe_sub(t, t, q); fe_sub(t, t, q);      // X3
    if (need_y) { fe_sub(q, q, t); fe_mul(q, r, q); fe_mul(u2, Y, ppp); fe_sub(Y, q, u2); fe_mul(ZZZ, ZZZ, ppp); }
    X = t;
    fe_mul(ZZ, ZZ, pp);
    return true;
}
// the same for an affine accumulator (ZZ = ZZZ = 1): 4M + 2S
VG_DEV bool vg_xyzz_add_affine_first(fe &X, fe &Y, fe &ZZ, fe &ZZZ, const fe &bx, const fe &by) {
    fe p, r, q, t, u;
    fe_sub(p, bx, X); fe_sub(r, by, Y);
    fe pc = p; fe_canon(pc);
    if (fe_is_zero_raw(pc)) return false;
    fe_sqr_fast(ZZ, p); fe_mul(ZZZ, ZZ, p); fe_mul(q, X, ZZ);
    fe_sqr_fast(t, r); fe_sub(t, t, ZZZ); fe_sub(t, t, q); fe_sub(t, t, q);      // X3
    fe_sub(q, q, t); fe_mul(q, r, q); fe_mul(u, Y, ZZZ); fe_sub(Y, q, u);
    X = t;
    return true;
}

// Q = P + t*G (VG_TAP_SCALAR: Q = t*G, P unused, t the full scalar) as qx = X and qzz = Z^2 of a projective representative
// (x(Q) = qx / qzz). Returns false when Q is the point at infinity.
// FAST = true: generic-case additions only; returns false at the first special case (the caller falls back to FAST = false).
// FAST = false: gej_add_affine handles every special case (equal points -> doubling, opposite points -> infinity).
template <bool FAST>
VG_DEV bool vg_tap_output_key(fe &qx, fe &qzz, const fe &px, const fe &py, const uint32_t t[8], const uint32_t *tab, uint32_t bits, bool active = true, uint32_t *viol = nullptr) {
    (void)viol;
    gej_t q;
#if VG_TAP_XYZZ
    fe zzz; fe_set_one(zzz);                    // fast path: q.z holds ZZ
#endif
#if VG_TAP_SCALAR
    (void)px; (void)py;
    fe_set_zero(q.x); fe_set_zero(q.y); fe_set_one(q.z); q.inf = true;
    bool special = false, affine = false, have = false;      // have: q holds a point (the first non-zero digit's table entry, as it is)
#else
    q.x = px; q.y = py; fe_set_one(q.z); q.inf = false;
    bool special = false, affine = true, have = true;        // affine: q.z is still 1
#endif
    const uint32_t n_win = vg_tap_windows(bits);
    const bool sgn = (bits & VG_TAP_SIGNED) != 0;
    (void)viol;
    uint32_t tt[8];
    VG_UNROLL for (int i = 0; i < 8; i++) tt[i] = t[i];
#if VG_TAP_SCALAR
    if (sgn) vg_tap_fold_scalar(tt);            // (the point form cannot flip the sign of t: signed geometries need the scalar form)
#endif
    uint32_t carry = 0;
#if VG_TAP_PREFETCH && !defined(VG_HOST_EMU)
    if (FAST && active) {
        #pragma unroll 1
        for (uint32_t j = 0; j < n_win; j++) {
            bool ng;
            const unsigned long long idx = vg_tap_entry(bits, tt, j, carry, ng);
            if (idx) {
                const uint32_t *e = tab + (size_t)idx * 16;
                asm volatile("prefetch.global.L2 [%0];" :: "l"(e));
                asm volatile("prefetch.global.L2 [%0];" :: "l"(e + 8));
            }
        }
        carry = 0;
    }
#endif
#ifndef VG_HOST_EMU
    #pragma unroll 1
#endif
    for (uint32_t j = 0; j < n_win; j++) {
#if VG_TAP_SYNC >= 2 && !defined(VG_HOST_EMU)
        if (FAST) __syncthreads();      // every thread of the CTA runs all n_win iterations (inactive ones skip the addition)
#endif
        bool neg;
        const unsigned long long idx = vg_tap_entry(bits, tt, j, carry, neg);
        if (idx == 0 || !active || special) continue;
        VG_ASSERT_AT(viol, j < n_win && idx < vg_tap_entries(bits), 10);
        const uint32_t *e = tab + (size_t)idx * 16;
        fe bx, by;
#ifdef VG_HOST_EMU
        for (int l = 0; l < 8; l++) { bx.v[l] = e[l]; by.v[l] = e[8 + l]; }
#else
        const uint4 *e4 = (const uint4 *)e;
        uint4 a0 = __ldg(e4), a1 = __ldg(e4 + 1), a2 = __ldg(e4 + 2), a3 = __ldg(e4 + 3);
        bx.v[0] = a0.x; bx.v[1] = a0.y; bx.v[2] = a0.z; bx.v[3] = a0.w; bx.v[4] = a1.x; bx.v[5] = a1.y; bx.v[6] = a1.z; bx.v[7] = a1.w;
        by.v[0] = a2.x; by.v[1] = a2.y; by.v[2] = a2.z; by.v[3] = a2.w; by.v[4] = a3.x; by.v[5] = a3.y; by.v[6] = a3.z; by.v[7] = a3.w;
#endif
        if (neg) { fe ny; fe_neg(ny, by); by = ny; }
        if (!have) { q.x = bx; q.y = by; fe_set_one(q.z); q.inf = false; have = true; affine = true; continue; }
        if (FAST) {
#if VG_TAP_XYZZ
            if (affine) { if (!vg_xyzz_add_affine_first(q.x, q.y, q.z, zzz, bx, by)) special = true; affine = false; }
            else if (!vg_xyzz_add_affine_fast(q.x, q.y, q.z, zzz, bx, by, !VG_TAP_LAST_NO_Y || j + 1 < n_win)) special = true;
#else
            if (VG_TAP_FIRST_AFFINE && affine) { if (!vg_aff_add_affine_fast(q, bx, by)) special = true; affine = false; }
            else if (!vg_gej_add_affine_fast(q, bx, by, !VG_TAP_LAST_NO_Y || j + 1 < n_win)) special = true;
#endif
        }
        else { gej_t r; gej_add_affine(r, q, bx, by); q = r; if (q.inf) { fe_set_one(q.z); } }
    }
    qx = q.x;
    if (FAST && VG_TAP_XYZZ) qzz = q.z; else fe_sqr_fast(qzz, q.z);
    return have && !q.inf && !special;
}

VG_COLD bool vg_tap_output_key_slow(fe &qx, fe &qzz, const fe &px, const fe &py, const uint32_t t[8], const uint32_t *tab, uint32_t bits) {
    return vg_tap_output_key<false>(qx, qzz, px, py, t, tab, bits);
}

// stage-2 thread: keys [first, first + VG_TAP_GROUP) of the window
template <int MODE>
VG_DEV void vg_tap_thread(const vg_scan_params &P, const vg_dfa_view &dv, unsigned long long first) {
    // X_k, Z_k^2 and the running product Z_0^2 ... Z_k^2 (local memory: indexed by the loop counter)
    fe X[VG_TAP_GROUP], Z2[VG_TAP_GROUP], pre[VG_TAP_GROUP];
    unsigned long long ok_mask = 0;
    fe run; fe_set_one(run);
#ifndef VG_HOST_EMU
    #pragma unroll 1
#endif
    for (int k = 0; k < VG_TAP_GROUP; k++) {
#if VG_TAP_SYNC == 1 && !defined(VG_HOST_EMU)
        __syncthreads();      // re-align the CTA's warps once per key (instruction-fetch sharing, see VG_SYNC_ITER in scan_core.cuh)
#endif
        VG_ASSERT(P, first + (unsigned long long)k < (unsigned long long)P.n_threads * 2ull * P.half, 6);
        const uint32_t *p = P.points + (first + (unsigned long long)k) * 16;
        fe px, py;
#ifdef VG_HOST_EMU
        for (int l = 0; l < 8; l++) { px.v[l] = p[l]; py.v[l] = p[8 + l]; }
#else
        const uint4 *p4 = (const uint4 *)p;
        uint4 a0 = p4[0], a1 = p4[1], a2 = p4[2], a3 = p4[3];
        px.v[0] = a0.x; px.v[1] = a0.y; px.v[2] = a0.z; px.v[3] = a0.w; px.v[4] = a1.x; px.v[5] = a1.y; px.v[6] = a1.z; px.v[7] = a1.w;
        py.v[0] = a2.x; py.v[1] = a2.y; py.v[2] = a2.z; py.v[3] = a2.w; py.v[4] = a3.x; py.v[5] = a3.y; py.v[6] = a3.z; py.v[7] = a3.w;
#endif
        bool ok = !(fe_is_zero_raw(px) && fe_is_zero_raw(py));
        fe z2; fe_set_one(z2);
        fe_set_zero(X[k]);
        {
            uint32_t t[8];
            vg_taptweak_hash(t, px);
            ok = ok && vg_tweak_is_valid(t);
#if VG_TAP_SCALAR
            {   // t <- s = t + k (even y) or t + n - k (odd y) mod n, k = window base + key index; s == 0: Q would be infinity
                uint32_t kk[8], tl[8], sl[8], any = 0;
                scalar_add_u64(kk, P.tap_base, first + (unsigned long long)k);
                if (py.v[0] & 1u) vg_scalar_neg_n(kk, kk);
                VG_UNROLL for (int i = 0; i < 8; i++) tl[i] = t[7 - i];
                vg_scalar_add_mod_n(sl, tl, kk);
                VG_UNROLL for (int i = 0; i < 8; i++) { t[7 - i] = sl[i]; any |= sl[i]; }
                ok = ok && any != 0;
            }
#endif
            fe qx, qzz;
            // all threads walk the table loop together (barriers inside); invalid keys skip the additions
            if (!vg_tap_output_key<true>(qx, qzz, px, py, t, P.tap_table, P.tap_bits, ok, P.violations) && ok)
                ok = vg_tap_output_key_slow(qx, qzz, px, py, t, P.tap_table, P.tap_bits);      // special case met (practically never)
            if (ok) { X[k] = qx; z2 = qzz; }
        }
        if (ok) ok_mask |= 1ull << k;
        Z2[k] = z2;
        fe_mul(run, run, z2);
        pre[k] = run;
    }
    fe inv;
    fe_inv(inv, run);                                   // 1 / (Z_0^2 ... Z_{G-1}^2)
#ifndef VG_HOST_EMU
    #pragma unroll 1
#endif
    for (int k = VG_TAP_GROUP - 1; k >= 0; k--) {
        fe zi2, xq;
        if (k > 0) { fe_mul(zi2, inv, pre[k - 1]); fe_mul(inv, inv, Z2[k]); }   // 1/Z_k^2, then inv -> 1/pre[k-1]
        else zi2 = inv;
        fe_mul(xq, X[k], zi2);
        fe_canon(xq);
        uint32_t dg[VG_DIGEST_WORDS];
        VG_UNROLL for (int i = 0; i < 8; i++) dg[i] = vg_bswap(xq.v[7 - i]);
        vg_finish_point<VG_FMT_P2TR, MODE>(P, dv, dg, first + (unsigned long long)k, ((ok_mask >> k) & 1ull) != 0);
    }
}
