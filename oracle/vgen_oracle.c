/* ORACLE — CPU restatement of vgen's deterministic CPU scan path. TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product (vgen_b200/, libvgen_cuda.so) never links or calls it.
 *
 * Follows:  scan_range_cpu            /root/reference/src/scanner.rs:211-330
 *           AddressGenerator::generate /root/reference/src/address.rs:92-151
 *           to_checksum_address        /root/reference/src/address.rs:176-198
 *           Pattern::new / matches     /root/reference/src/pattern.rs:21-45
 * The arithmetic itself lives in un-vendored crates (bitcoin 0.32.8, secp256k1-sys 0.10.1,
 * bitcoin_hashes 0.14.1, base58ck 0.1.0, bech32 0.11.1, sha3 0.10.8, regex 1.12.2 — Cargo.lock),
 * so it is restated from the published standards (SEC2, FIPS 180-4, RIPEMD-160, Keccak, BIP173/350,
 * BIP341, EIP-55). Parity is PINNED by: the reference's own known answers (address.rs:233-238,
 * provider.rs:76-87, README.md:107-108,121-127) and by Python `cryptography`/`hashlib` vectors in
 * tests/golden/ (tests/golden/make_golden.py). The reference itself (Rust) cannot be built here.
 *
 * Algorithm class for k*G: fixed-base windowed table (8-bit windows), one independent scalar
 * multiplication per key — as libsecp256k1's ecmult_gen does and as scan_range_cpu requires
 * (scanner.rs:287-294: every key goes through SecretKey::from_slice + a full multiplication). */
#include <pthread.h>
#include <stdio.h>
#include <time.h>
#include "fe.h"
#include "hashes.h"
#include "encode.h"
#include "regex.h"

enum { FMT_P2PKH = 0, FMT_P2PKH_UNCOMPRESSED = 1, FMT_P2WPKH = 2, FMT_P2SH_P2WPKH = 3, FMT_P2TR = 4, FMT_ETHEREUM = 5 };

static ge G_TABLE[32][256];          /* G_TABLE[w][j] = j * 2^(8w) * G */
static int g_table_ready = 0;
static pthread_mutex_t g_init_lock = PTHREAD_MUTEX_INITIALIZER;

static const uint8_t GX_BE[32] = {0x79,0xBE,0x66,0x7E,0xF9,0xDC,0xBB,0xAC,0x55,0xA0,0x62,0x95,0xCE,0x87,0x0B,0x07,
                                  0x02,0x9B,0xFC,0xDB,0x2D,0xCE,0x28,0xD9,0x59,0xF2,0x81,0x5B,0x16,0xF8,0x17,0x98};
static const uint8_t GY_BE[32] = {0x48,0x3A,0xDA,0x77,0x26,0xA3,0xC4,0x65,0x5D,0xA4,0xFB,0xFC,0x0E,0x11,0x08,0xA8,
                                  0xFD,0x17,0xB4,0x48,0xA6,0x85,0x54,0x19,0x9C,0x47,0xD0,0x8F,0xFB,0x10,0xD4,0xB8};

int vgo_init(void) {
    pthread_mutex_lock(&g_init_lock);
    if (!g_table_ready) {
        ge base; fe_from_be(&base.x, GX_BE); fe_from_be(&base.y, GY_BE); base.inf = 0;
        for (int w = 0; w < 32; w++) {
            gej acc; acc.inf = 1;
            G_TABLE[w][0].inf = 1; fe_set_u64(&G_TABLE[w][0].x, 0); fe_set_u64(&G_TABLE[w][0].y, 0);
            for (int j = 1; j < 256; j++) { gej_add_ge(&acc, &acc, &base); ge_set_gej(&G_TABLE[w][j], &acc); }
            gej_add_ge(&acc, &acc, &base);              /* 256 * base */
            ge_set_gej(&base, &acc);
        }
        g_table_ready = 1;
    }
    pthread_mutex_unlock(&g_init_lock);
    return 0;
}
static int scalar_valid(const uint64_t k[4]) { return !u256_is_zero(k) && u256_cmp(k, SC_N) < 0; }

static void ecmult_gen(ge *r, const uint64_t k[4]) {
    gej acc; acc.inf = 1;
    for (int w = 0; w < 32; w++) {
        int byte = (int)((k[w >> 3] >> (8 * (w & 7))) & 0xff);
        if (byte) gej_add_ge(&acc, &acc, &G_TABLE[w][byte]);
    }
    ge_set_gej(r, &acc);
}
/* public key of a 32-byte big-endian secret; -1 if the secret is 0 or >= n (address.rs:93) */
int vgo_pubkey(const uint8_t key[32], uint8_t x[32], uint8_t y[32]) {
    uint64_t k[4]; u256_from_be(k, key);
    if (!scalar_valid(k)) return -1;
    if (!g_table_ready) vgo_init();
    ge p; ecmult_gen(&p, k);
    fe_to_be(x, &p.x); fe_to_be(y, &p.y);
    return 0;
}
static void tagged_hash_taptweak(const uint8_t x[32], uint8_t out[32]) {
    uint8_t tag[32], buf[96];
    sha256((const uint8_t *)"TapTweak", 8, tag);
    memcpy(buf, tag, 32); memcpy(buf + 32, tag, 32); memcpy(buf + 64, x, 32);
    sha256(buf, 96, out);
}
/* digest for a format from an affine public key; returns digest length (20 or 32) or -1 */
static int format_digest(int format, const ge *p, uint8_t digest[32]) {
    uint8_t pub[65];
    fe_to_be(pub + 1, &p->x);
    switch (format) {
        case FMT_P2PKH: case FMT_P2WPKH:
            pub[0] = fe_is_odd(&p->y) ? 3 : 2; hash160(pub, 33, digest); return 20;
        case FMT_P2PKH_UNCOMPRESSED:
            pub[0] = 4; fe_to_be(pub + 33, &p->y); hash160(pub, 65, digest); return 20;
        case FMT_P2SH_P2WPKH: {
            uint8_t script[22]; script[0] = 0x00; script[1] = 0x14;
            pub[0] = fe_is_odd(&p->y) ? 3 : 2; hash160(pub, 33, script + 2);
            hash160(script, 22, digest); return 20;
        }
        case FMT_ETHEREUM: {
            uint8_t h[32]; fe_to_be(pub + 33, &p->y); keccak256(pub + 1, 64, h);
            memcpy(digest, h + 12, 20); return 20;
        }
        case FMT_P2TR: {
            /* BIP341 key-path output with no script tree: Q = lift_x(x(P)) + H_TapTweak(x(P))*G */
            uint8_t t[32]; uint64_t tk[4];
            tagged_hash_taptweak(pub + 1, t); u256_from_be(tk, t);
            if (u256_cmp(tk, SC_N) >= 0) return -1;
            ge pe = *p; if (fe_is_odd(&pe.y)) fe_neg(&pe.y, &pe.y);
            gej q; gej_set_ge(&q, &pe);
            if (!u256_is_zero(tk)) { ge tg; ecmult_gen(&tg, tk); gej_add_ge(&q, &q, &tg); }
            ge qa; ge_set_gej(&qa, &q);
            if (qa.inf) return -1;
            fe_to_be(digest, &qa.x); return 32;
        }
    }
    return -1;
}
static int format_address(int format, const uint8_t *digest, char *addr) {
    uint8_t payload[21];
    switch (format) {
        case FMT_P2PKH: case FMT_P2PKH_UNCOMPRESSED:
            payload[0] = 0x00; memcpy(payload + 1, digest, 20); return base58check_encode(payload, 21, addr);
        case FMT_P2SH_P2WPKH:
            payload[0] = 0x05; memcpy(payload + 1, digest, 20); return base58check_encode(payload, 21, addr);
        case FMT_P2WPKH: return segwit_encode("bc", 0, digest, 20, addr);
        case FMT_P2TR: return segwit_encode("bc", 1, digest, 32, addr);
        case FMT_ETHEREUM: eip55_encode(digest, addr); return 42;
    }
    return -1;
}
/* address string from a digest alone (what the reference GPU loop does on the host, gpu.rs:1034-1060) */
int vgo_address_from_digest(int format, const uint8_t *digest, char *addr) { return format_address(format, digest, addr); }

/* AddressGenerator::generate: address, WIF, digest. Returns 0 or -1 for an invalid secret. */
int vgo_address(int format, const uint8_t key[32], char *addr, char *wif, uint8_t digest[32]) {
    uint64_t k[4]; u256_from_be(k, key);
    if (!scalar_valid(k)) return -1;
    if (!g_table_ready) vgo_init();
    ge p; ecmult_gen(&p, k);
    uint8_t dg[32]; memset(dg, 0, 32);
    if (format_digest(format, &p, dg) < 0) return -1;
    if (digest) memcpy(digest, dg, 32);
    if (addr) format_address(format, dg, addr);
    if (wif) {
        if (format == FMT_ETHEREUM) hex_encode(key, 32, wif);          /* address.rs:110 */
        else wif_encode(key, format != FMT_P2PKH_UNCOMPRESSED, wif);  /* address.rs:116-118,146 */
    }
    return 0;
}
/* hash160 of the compressed key for n sequential scalars (the reference GPU contract, gpu.rs:660) */
int vgo_hash160_batch(const uint8_t start[32], uint64_t n, uint8_t *out20) {
    uint64_t k[4]; u256_from_be(k, start);
    if (!g_table_ready) vgo_init();
    for (uint64_t i = 0; i < n; i++) {
        if (scalar_valid(k)) { ge p; ecmult_gen(&p, k); uint8_t d[32]; format_digest(FMT_P2PKH, &p, d); memcpy(out20 + 20 * i, d, 20); }
        else memset(out20 + 20 * i, 0, 20);
        u256_add_u64(k, k, 1, NULL);
    }
    return 0;
}
int vgo_digest_batch(int format, const uint8_t start[32], uint64_t n, uint8_t *out32) {
    uint64_t k[4]; u256_from_be(k, start);
    if (!g_table_ready) vgo_init();
    for (uint64_t i = 0; i < n; i++) {
        memset(out32 + 32 * i, 0, 32);
        if (scalar_valid(k)) { ge p; ecmult_gen(&p, k); format_digest(format, &p, out32 + 32 * i); }
        u256_add_u64(k, k, 1, NULL);
    }
    return 0;
}

/* ---- thin wrappers so tests can pin each primitive ---- */
void vgo_sha256(const uint8_t *m, uint64_t n, uint8_t *o) { sha256(m, n, o); }
void vgo_ripemd160(const uint8_t *m, uint64_t n, uint8_t *o) { ripemd160(m, n, o); }
void vgo_keccak256(const uint8_t *m, uint64_t n, uint8_t *o) { keccak256(m, n, o); }
int vgo_base58check(const uint8_t *m, int n, char *o) { return base58check_encode(m, n, o); }
int vgo_base58_decode(const char *s, uint8_t *o, int cap) { return base58_decode(s, o, cap); }
int vgo_segwit(int witver, const uint8_t *prog, int n, char *o) { return segwit_encode("bc", witver, prog, n, o); }
int vgo_base58(const uint8_t *m, int n, char *o) { return base58_encode(m, n, o); }
/* BIP341 key-path output key (no script tree) of an x-only INTERNAL key: lift_x (even y = (x^3 + 7)^((p+1)/4)), then the same
 * tweak as format_digest(FMT_P2TR). For the published BIP341 / BIP86 vectors, which give internal public keys. -1: x not on the curve. */
int vgo_taproot_output(const uint8_t xonly[32], uint8_t out[32]) {
    static const uint64_t E[4] = {0xFFFFFFFFBFFFFF0CULL, ~0ULL, ~0ULL, 0x3FFFFFFFFFFFFFFFULL};   /* (p + 1) / 4 */
    ge p; fe rhs, y, t, seven;
    if (!g_table_ready) vgo_init();
    fe_from_be(&p.x, xonly);
    if (u256_cmp(p.x.d, FE_P) >= 0) return -1;
    fe_sqr(&t, &p.x); fe_mul(&rhs, &t, &p.x); fe_set_u64(&seven, 7); fe_add(&rhs, &rhs, &seven);
    fe_set_u64(&y, 1);
    for (int i = 255; i >= 0; i--) { fe_sqr(&y, &y); if ((E[i >> 6] >> (i & 63)) & 1) fe_mul(&y, &y, &rhs); }
    fe_sqr(&t, &y);
    if (!fe_eq(&t, &rhs)) return -1;
    if (fe_is_odd(&y)) fe_neg(&y, &y);
    p.y = y; p.inf = 0;
    return format_digest(FMT_P2TR, &p, out) == 32 ? 0 : -1;
}
void vgo_eip55(const uint8_t *a, char *o) { eip55_encode(a, o); }
int vgo_fe_selftest(void) {
    fe a, b, c, one; fe_from_be(&a, GX_BE); fe_inv(&b, &a); fe_mul(&c, &a, &b); fe_set_u64(&one, 1);
    return fe_eq(&c, &one) ? 0 : -1;
}
int vgo_regex_check(const char *pattern, int ci, char *err, int errcap) {
    oregex re; int rc = oregex_compile(&re, pattern, ci);
    if (rc && err) snprintf(err, errcap, "%s", re.err);
    oregex_free(&re); return rc;
}
int vgo_regex_match(const char *pattern, int ci, const char *s) {
    oregex re; if (oregex_compile(&re, pattern, ci)) return -1;
    int r = oregex_match(&re, s, strlen(s)); oregex_free(&re); return r;
}

/* ---- scan_range_cpu (scanner.rs:211-330) ---- */
typedef struct {
    int format; const oregex *re; uint64_t start[4]; uint64_t n_keys, n_batches;
    uint64_t next_batch; pthread_mutex_t lock;
    uint64_t ops; uint8_t *keys_out; uint64_t cap, found;
} scan_job;
#define SCAN_BATCH 10000ULL      /* scanner.rs:241 */

static void *scan_worker(void *arg) {
    scan_job *job = (scan_job *)arg;
    uint8_t (*local)[32] = (uint8_t (*)[32])malloc(32 * SCAN_BATCH);
    for (;;) {
        pthread_mutex_lock(&job->lock);
        uint64_t b = job->next_batch++;
        pthread_mutex_unlock(&job->lock);
        if (b >= job->n_batches) break;
        uint64_t off = b * SCAN_BATCH, cnt = (b == job->n_batches - 1) ? job->n_keys - off : SCAN_BATCH;
        uint64_t k[4]; u256_add_u64(k, job->start, off, NULL);
        uint64_t ops = 0, nf = 0;
        for (uint64_t i = 0; i < cnt; i++) {
            if (scalar_valid(k)) {                                  /* scanner.rs:294: None -> skipped, not counted */
                ge p; ecmult_gen(&p, k);
                uint8_t dg[32]; char addr[96];
                if (format_digest(job->format, &p, dg) >= 0) {
                    ops++;
                    int n = format_address(job->format, dg, addr);
                    if (oregex_match(job->re, addr, (size_t)n)) u256_to_be(local[nf++], k);
                }
            }
            u256_add_u64(k, k, 1, NULL);
        }
        pthread_mutex_lock(&job->lock);
        job->ops += ops;
        for (uint64_t i = 0; i < nf; i++) {
            if (job->found < job->cap) memcpy(job->keys_out + 32 * job->found, local[i], 32);
            job->found++;
        }
        pthread_mutex_unlock(&job->lock);
    }
    free(local);
    return NULL;
}
/* Scans the inclusive window [start, start+n_keys-1]; writes matching secrets (32 B big-endian each,
 * unordered like the reference) to keys_out up to cap; *found gets the total match count, *ops the
 * number of valid keys tested. Returns 0, or -1 on a bad pattern. */
int vgo_scan_range(int format, const char *pattern, int ci, const uint8_t start[32], uint64_t n_keys, int threads,
                   uint8_t *keys_out, uint64_t cap, uint64_t *found, uint64_t *ops, double *elapsed) {
    oregex re; if (oregex_compile(&re, pattern, ci)) return -1;
    vgo_init();
    scan_job job; memset(&job, 0, sizeof job);
    job.format = format; job.re = &re; u256_from_be(job.start, start);
    job.n_keys = n_keys; job.n_batches = (n_keys + SCAN_BATCH - 1) / SCAN_BATCH;
    job.keys_out = keys_out; job.cap = cap; pthread_mutex_init(&job.lock, NULL);
    if (threads < 1) threads = 1;
    if (threads > 256) threads = 256;
    struct timespec t0, t1; clock_gettime(CLOCK_MONOTONIC, &t0);
    pthread_t th[256];
    for (int i = 0; i < threads; i++) pthread_create(&th[i], NULL, scan_worker, &job);
    for (int i = 0; i < threads; i++) pthread_join(th[i], NULL);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    if (found) *found = job.found;
    if (ops) *ops = job.ops;
    if (elapsed) *elapsed = (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
    oregex_free(&re); pthread_mutex_destroy(&job.lock);
    return 0;
}
