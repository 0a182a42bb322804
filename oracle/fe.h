/* ORACLE (test infrastructure, not product): secp256k1 field + curve arithmetic on the CPU.
 *
 * Restates the arithmetic the reference reaches through bitcoin 0.32.8 -> secp256k1 0.29.1 ->
 * secp256k1-sys 0.10.1 (Cargo.lock:144,1607,1618; call site src/address.rs:93-142), from the
 * published SEC2 curve parameters. Only tests/, __graft_entry__.smoke() and bench.py's CPU
 * baseline may use anything under oracle/.
 */
#ifndef VGEN_ORACLE_FE_H
#define VGEN_ORACLE_FE_H
#include <stdint.h>
#include <string.h>

typedef unsigned __int128 u128;
typedef struct { uint64_t d[4]; } fe;            /* little-endian 64-bit limbs, always < p */
typedef struct { fe x, y; int inf; } ge;         /* affine point */
typedef struct { fe x, y, z; int inf; } gej;     /* Jacobian point */

#define FE_RC 0x1000003D1ULL                     /* 2^256 mod p */
static const uint64_t FE_P[4] = {0xFFFFFFFEFFFFFC2FULL, ~0ULL, ~0ULL, ~0ULL};
static const uint64_t SC_N[4] = {0xBFD25E8CD0364141ULL, 0xBAAEDCE6AF48A03BULL,
                                 0xFFFFFFFFFFFFFFFEULL, 0xFFFFFFFFFFFFFFFFULL};

static inline int u256_cmp(const uint64_t a[4], const uint64_t b[4]) {
    for (int i = 3; i >= 0; i--) { if (a[i] < b[i]) return -1; if (a[i] > b[i]) return 1; }
    return 0;
}
static inline int u256_is_zero(const uint64_t a[4]) { return (a[0] | a[1] | a[2] | a[3]) == 0; }
static inline uint64_t u256_add(uint64_t r[4], const uint64_t a[4], const uint64_t b[4]) {
    u128 c = 0;
    for (int i = 0; i < 4; i++) { c += (u128)a[i] + b[i]; r[i] = (uint64_t)c; c >>= 64; }
    return (uint64_t)c;
}
static inline uint64_t u256_sub(uint64_t r[4], const uint64_t a[4], const uint64_t b[4]) {
    uint64_t br = 0;
    for (int i = 0; i < 4; i++) {
        u128 t = (u128)a[i] - b[i] - br; r[i] = (uint64_t)t; br = (uint64_t)(t >> 64) & 1;
    }
    return br;
}
static inline void u256_add_u64(uint64_t r[4], const uint64_t a[4], uint64_t v, uint64_t *carry) {
    u128 c = v;
    for (int i = 0; i < 4; i++) { c += a[i]; r[i] = (uint64_t)c; c >>= 64; }
    if (carry) *carry = (uint64_t)c;
}
static inline void u256_from_be(uint64_t r[4], const uint8_t b[32]) {
    for (int i = 0; i < 4; i++) {
        uint64_t v = 0;
        for (int j = 0; j < 8; j++) v = (v << 8) | b[(3 - i) * 8 + j];
        r[i] = v;
    }
}
static inline void u256_to_be(uint8_t b[32], const uint64_t a[4]) {
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 8; j++) b[(3 - i) * 8 + j] = (uint8_t)(a[i] >> (56 - 8 * j));
}

static inline void fe_set_u64(fe *r, uint64_t v) { r->d[0] = v; r->d[1] = r->d[2] = r->d[3] = 0; }
static inline int fe_is_zero(const fe *a) { return u256_is_zero(a->d); }
static inline int fe_eq(const fe *a, const fe *b) { return u256_cmp(a->d, b->d) == 0; }
static inline int fe_is_odd(const fe *a) { return (int)(a->d[0] & 1); }

static inline void fe_add(fe *r, const fe *a, const fe *b) {
    uint64_t c = u256_add(r->d, a->d, b->d);
    if (c || u256_cmp(r->d, FE_P) >= 0) u256_sub(r->d, r->d, FE_P);
}
static inline void fe_sub(fe *r, const fe *a, const fe *b) {
    if (u256_sub(r->d, a->d, b->d)) u256_add(r->d, r->d, FE_P);
}
static inline void fe_neg(fe *r, const fe *a) {
    fe z; fe_set_u64(&z, 0); fe_sub(r, &z, a);
}
/* r = t mod p for a 512-bit t: fold the high half by 2^256 = FE_RC (mod p), twice. */
static inline void fe_reduce512(fe *r, const uint64_t t[8]) {
    u128 acc = 0; uint64_t lo[4];
    for (int i = 0; i < 4; i++) {
        acc += (u128)t[i] + (u128)t[4 + i] * FE_RC; lo[i] = (uint64_t)acc; acc >>= 64;
    }
    uint64_t c = (uint64_t)acc;                    /* < 2^34 */
    acc = (u128)lo[0] + (u128)c * FE_RC; lo[0] = (uint64_t)acc; acc >>= 64;
    for (int i = 1; i < 4; i++) { acc += lo[i]; lo[i] = (uint64_t)acc; acc >>= 64; }
    if ((uint64_t)acc) {                           /* wrapped once more: value is tiny, no further carry */
        acc = (u128)lo[0] + FE_RC; lo[0] = (uint64_t)acc; acc >>= 64;
        for (int i = 1; i < 4; i++) { acc += lo[i]; lo[i] = (uint64_t)acc; acc >>= 64; }
    }
    memcpy(r->d, lo, sizeof lo);
    if (u256_cmp(r->d, FE_P) >= 0) u256_sub(r->d, r->d, FE_P);
}
static inline void fe_mul(fe *r, const fe *a, const fe *b) {
    uint64_t t[8] = {0};
    for (int i = 0; i < 4; i++) {
        u128 c = 0;
        for (int j = 0; j < 4; j++) {
            c += (u128)a->d[i] * b->d[j] + t[i + j]; t[i + j] = (uint64_t)c; c >>= 64;
        }
        t[i + 4] = (uint64_t)c;
    }
    fe_reduce512(r, t);
}
static inline void fe_sqr(fe *r, const fe *a) { fe_mul(r, a, a); }
static inline void fe_sqr_n(fe *r, const fe *a, int n) {
    *r = *a; for (int i = 0; i < n; i++) fe_sqr(r, r);
}
/* a^(p-2): p-2 = 223 ones, 0, 22 ones, 0000, 1, 0, 11, 0, 1 (binary, msb first). */
static inline void fe_inv(fe *r, const fe *a) {
    fe x2, x3, x6, x9, x11, x22, x44, x88, x176, x220, x223, t;
    fe_sqr(&t, a); fe_mul(&x2, &t, a);
    fe_sqr(&t, &x2); fe_mul(&x3, &t, a);
    fe_sqr_n(&t, &x3, 3); fe_mul(&x6, &t, &x3);
    fe_sqr_n(&t, &x6, 3); fe_mul(&x9, &t, &x3);
    fe_sqr_n(&t, &x9, 2); fe_mul(&x11, &t, &x2);
    fe_sqr_n(&t, &x11, 11); fe_mul(&x22, &t, &x11);
    fe_sqr_n(&t, &x22, 22); fe_mul(&x44, &t, &x22);
    fe_sqr_n(&t, &x44, 44); fe_mul(&x88, &t, &x44);
    fe_sqr_n(&t, &x88, 88); fe_mul(&x176, &t, &x88);
    fe_sqr_n(&t, &x176, 44); fe_mul(&x220, &t, &x44);
    fe_sqr_n(&t, &x220, 3); fe_mul(&x223, &t, &x3);
    fe_sqr_n(&t, &x223, 23); fe_mul(&t, &t, &x22);
    fe_sqr_n(&t, &t, 5); fe_mul(&t, &t, a);
    fe_sqr_n(&t, &t, 3); fe_mul(&t, &t, &x2);
    fe_sqr_n(&t, &t, 2); fe_mul(r, &t, a);
}
static inline void fe_from_be(fe *r, const uint8_t b[32]) { u256_from_be(r->d, b); }
static inline void fe_to_be(uint8_t b[32], const fe *a) { u256_to_be(b, a->d); }

/* ---- curve: y^2 = x^3 + 7 ---- */
static inline void gej_set_ge(gej *r, const ge *a) {
    r->x = a->x; r->y = a->y; fe_set_u64(&r->z, 1); r->inf = a->inf;
}
static inline void gej_double(gej *r, const gej *a) {
    if (a->inf || fe_is_zero(&a->y)) { r->inf = 1; return; }
    fe A, B, C, D, E, F, t, x3, y3, z3;
    fe_sqr(&A, &a->x); fe_sqr(&B, &a->y); fe_sqr(&C, &B);
    fe_add(&t, &a->x, &B); fe_sqr(&t, &t); fe_sub(&t, &t, &A); fe_sub(&t, &t, &C);
    fe_add(&D, &t, &t);                                   /* D = 2((X+B)^2 - A - C) */
    fe_add(&E, &A, &A); fe_add(&E, &E, &A);               /* E = 3A */
    fe_sqr(&F, &E);
    fe_sub(&x3, &F, &D); fe_sub(&x3, &x3, &D);
    fe_sub(&t, &D, &x3); fe_mul(&y3, &E, &t);
    fe_add(&t, &C, &C); fe_add(&t, &t, &t); fe_add(&t, &t, &t);   /* 8C */
    fe_sub(&y3, &y3, &t);
    fe_mul(&z3, &a->y, &a->z); fe_add(&z3, &z3, &z3);
    r->x = x3; r->y = y3; r->z = z3; r->inf = 0;
}
/* r = a + b with b affine */
static inline void gej_add_ge(gej *r, const gej *a, const ge *b) {
    if (b->inf) { *r = *a; return; }
    if (a->inf) { gej_set_ge(r, b); return; }
    fe z2, u2, s2, h, rr, h2, h3, v, t, x3, y3, z3;
    fe_sqr(&z2, &a->z); fe_mul(&u2, &b->x, &z2);
    fe_mul(&s2, &b->y, &z2); fe_mul(&s2, &s2, &a->z);
    fe_sub(&h, &u2, &a->x); fe_sub(&rr, &s2, &a->y);
    if (fe_is_zero(&h)) {
        if (fe_is_zero(&rr)) { gej_double(r, a); return; }
        r->inf = 1; return;
    }
    fe_sqr(&h2, &h); fe_mul(&h3, &h2, &h); fe_mul(&v, &a->x, &h2);
    fe_sqr(&x3, &rr); fe_sub(&x3, &x3, &h3); fe_sub(&x3, &x3, &v); fe_sub(&x3, &x3, &v);
    fe_sub(&t, &v, &x3); fe_mul(&y3, &rr, &t); fe_mul(&t, &a->y, &h3); fe_sub(&y3, &y3, &t);
    fe_mul(&z3, &a->z, &h);
    r->x = x3; r->y = y3; r->z = z3; r->inf = 0;
}
static inline void ge_set_gej(ge *r, const gej *a) {
    if (a->inf) { r->inf = 1; fe_set_u64(&r->x, 0); fe_set_u64(&r->y, 0); return; }
    fe zi, zi2, zi3;
    fe_inv(&zi, &a->z); fe_sqr(&zi2, &zi); fe_mul(&zi3, &zi2, &zi);
    fe_mul(&r->x, &a->x, &zi2); fe_mul(&r->y, &a->y, &zi3); r->inf = 0;
}
#endif
