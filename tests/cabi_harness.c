/* A second, non-Python client of include/vgen_cuda.h: plain C99, compiled with `gcc -Iinclude` and linked against
 * libvgen_cuda.so only. It makes the call sequence the Rust binding (src/gpu.rs) makes:
 *
 *   create -> set_pattern -> scan (progress callback + stop flag, small caller buffer) -> scan_total_hits / scan_fetch_hits
 *   -> format_address / format_wif per hit -> two "frames" (hash160_batch, digest_batch with VGEN_P2TR_INTERNAL_KEY) -> destroy
 *
 * and prints what it got, one record per line, for tests/test_cabi_harness.py to compare with the oracle.
 *
 *   cabi_harness <format 0..5> <pattern> <start hex> <n keys> <max_hits> [stop_after_callbacks]
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "vgen_cuda.h"

struct hooks { volatile int32_t stop; int calls, stop_after; uint64_t last; int monotone; };

static void on_progress(uint64_t keys_scanned, void *user) {
    struct hooks *h = (struct hooks *)user;
    if (keys_scanned < h->last) h->monotone = 0;
    h->last = keys_scanned;
    h->calls++;
    if (h->stop_after > 0 && h->calls >= h->stop_after) h->stop = 1;
}

static void hex_to_be32(const char *hex, uint8_t out[32]) {
    size_t n = strlen(hex), i;
    memset(out, 0, 32);
    for (i = 0; i < n && i < 64; i++) {
        char c = hex[n - 1 - i];
        int v = (c >= '0' && c <= '9') ? c - '0' : (c >= 'a' && c <= 'f') ? c - 'a' + 10 : (c >= 'A' && c <= 'F') ? c - 'A' + 10 : 0;
        out[31 - i / 2] |= (uint8_t)(v << (4 * (i & 1)));
    }
}
static void add_u64(uint8_t k[32], uint64_t v) {
    int i;
    unsigned carry = 0;
    for (i = 31; i >= 0; i--) {
        unsigned add = (unsigned)(v & 0xff) + carry + k[i];
        k[i] = (uint8_t)add; carry = add >> 8; v >>= 8;
    }
}
static void print_hex(const uint8_t *p, size_t n) { size_t i; for (i = 0; i < n; i++) printf("%02x", p[i]); }

#define CHECK(call) do { int rc_ = (call); if (rc_ != VGEN_OK) { fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, vgen_cuda_last_error(ctx)); return 2; } } while (0)

int main(int argc, char **argv) {
    vgen_ctx *ctx = NULL;
    uint8_t start[32], end[32];
    vgen_hit small[4], *all = NULL;
    uint64_t n_hits = 0, scanned = 0, total, got, i, n_keys, max_hits;
    struct hooks h;
    int format, dev = 0;
    char name[256];
    if (argc < 6) { fprintf(stderr, "usage: %s <format> <pattern> <start hex> <n keys> <max_hits> [stop_after_callbacks]\n", argv[0]); return 1; }
    format = atoi(argv[1]);
    hex_to_be32(argv[3], start);
    n_keys = strtoull(argv[4], NULL, 10);
    max_hits = strtoull(argv[5], NULL, 10);
    memcpy(end, start, 32);
    add_u64(end, n_keys - 1);
    memset(&h, 0, sizeof h);
    h.monotone = 1;
    h.stop_after = argc > 6 ? atoi(argv[6]) : 0;

    printf("DEVICES %d\n", vgen_cuda_device_count());
    if (vgen_cuda_create(&dev, 1, 524288u, &ctx) != VGEN_OK) { fprintf(stderr, "create failed: %s\n", vgen_cuda_last_error(NULL)); return 2; }
    CHECK(vgen_cuda_device_name(0, name, sizeof name));
    printf("DEVICE %s\n", name);
    printf("WINDOW %llu LANES %d\n", (unsigned long long)vgen_cuda_window_keys(ctx), vgen_cuda_lane_count(ctx));

    /* an invalid pattern is an error code + message, never an abort */
    if (vgen_cuda_set_pattern(ctx, format, "(", 0) != VGEN_ERR_PATTERN) { fprintf(stderr, "bad pattern accepted\n"); return 3; }
    printf("BADPATTERN %s\n", vgen_cuda_last_error(ctx));
    CHECK(vgen_cuda_set_pattern(ctx, format, argv[2], 0));
    printf("MATCHER %s\n", vgen_cuda_matcher_kind(ctx));

    /* scan with a 4-record caller buffer: the rest is read back with scan_fetch_hits */
    CHECK(vgen_cuda_scan(ctx, start, end, max_hits, &h.stop, on_progress, &h, small, 4, &n_hits, &scanned));
    total = vgen_cuda_scan_total_hits(ctx);
    printf("SCANNED %llu TOTAL %llu FIRST %llu CALLBACKS %d MONOTONE %d LAST %llu STOPPED %d\n", (unsigned long long)scanned, (unsigned long long)total,
           (unsigned long long)n_hits, h.calls, h.monotone, (unsigned long long)h.last, (int)h.stop);
    all = (vgen_hit *)calloc(total ? total : 1, sizeof(vgen_hit));
    memcpy(all, small, (size_t)n_hits * sizeof(vgen_hit));
    for (got = n_hits; got < total;) {
        uint64_t m = vgen_cuda_scan_fetch_hits(ctx, got, all + got, total - got);
        if (!m) break;
        got += m;
    }
    for (i = 0; i < got; i++) {
        char addr[128], wif[128];
        if (vgen_format_address(format, all[i].digest, addr, sizeof addr) < 0 || vgen_format_wif(format, all[i].key_be, wif, sizeof wif) < 0) return 4;
        printf("HIT "); print_hex(all[i].key_be, 32); printf(" %s %s %u\n", addr, wif, all[i].digest_len);
    }
    free(all);

    /* the reference's frame API: two hash160 frames back to back, then a P2TR internal-key frame */
    {
        const uint64_t n = 4096;
        uint8_t *f0 = (uint8_t *)malloc(20 * n), *f1 = (uint8_t *)malloc(20 * n), *fx = (uint8_t *)malloc(32 * n);
        uint8_t second[32];
        memcpy(second, start, 32);
        add_u64(second, n);
        CHECK(vgen_cuda_hash160_batch(ctx, start, n, f0));
        CHECK(vgen_cuda_hash160_batch(ctx, second, n, f1));
        CHECK(vgen_cuda_digest_batch(ctx, VGEN_P2TR_INTERNAL_KEY, start, n, fx, 32));
        printf("FRAME0 "); print_hex(f0, 20 * n); printf("\n");
        printf("FRAME1 "); print_hex(f1, 20 * n); printf("\n");
        printf("FRAMEX "); print_hex(fx, 32 * 8); printf("\n");
        free(f0); free(f1); free(fx);
    }
    /* end < start is VGEN_ERR_RANGE */
    if (vgen_cuda_scan(ctx, end, start, 0, NULL, NULL, NULL, small, 4, &n_hits, &scanned) != VGEN_ERR_RANGE && n_keys > 1) { fprintf(stderr, "reversed range accepted\n"); return 5; }
    vgen_cuda_destroy(ctx);
    printf("DONE\n");
    return 0;
}
