/* libvgen_cuda — C ABI of the B200 (sm_100a) keygen-and-match engine.
 *
 * Drop-in boundary for vgen's GPU scan path: these entry points are what a new src/gpu.rs binds instead
 * of wgpu (see INTEGRATION.md for the Rust `extern "C"` block and the ctypes binding used by tests/).
 * Plain pointers and sizes only; no CUDA or torch types in any signature. All functions return
 * VGEN_OK (0) or a negative error class; vgen_cuda_last_error() gives the message. Nothing throws or
 * aborts across the boundary. One scan at a time per context.
 *
 * Reference interface replaced (file:line in oritwoen/vgen v0.3.0):
 *   GpuRunner::new                    src/gpu.rs:138-533   -> vgen_cuda_create
 *   GpuRunner::{device_name,backend}  src/gpu.rs:128-136   -> vgen_cuda_device_name
 *   GpuRunner::generate_batch         src/gpu.rs:660-665   -> vgen_cuda_hash160_batch
 *   GpuRunner::dispatch/await_result  src/gpu.rs:535-658   -> (internal; two windows in flight per GPU)
 *   scan_gpu_with_runner / scan_gpu   src/gpu.rs:920-1166  -> vgen_cuda_set_pattern + vgen_cuda_scan
 *   Pattern::new / Pattern::matches   src/pattern.rs:21-45 -> vgen_cuda_set_pattern / vgen_pattern_is_match
 *   hash160 -> address string, WIF    src/gpu.rs:1034-1060,1080; src/address.rs:92-151 -> vgen_format_address / vgen_format_wif
 */
#ifndef VGEN_CUDA_H
#define VGEN_CUDA_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VGEN_OK 0
#define VGEN_ERR_INVALID_ARG (-1)
#define VGEN_ERR_CUDA (-2)       /* a CUDA call failed, or no sm_100 device */
#define VGEN_ERR_PATTERN (-3)    /* empty / invalid / unsupported regex (src/pattern.rs:22-33) */
#define VGEN_ERR_RANGE (-4)      /* end < start */
#define VGEN_ERR_BUFFER (-5)     /* caller buffer too small */

/* AddressFormat (src/address.rs:11-46) */
enum vgen_format {
    VGEN_P2PKH = 0, VGEN_P2PKH_UNCOMPRESSED = 1, VGEN_P2WPKH = 2, VGEN_P2SH_P2WPKH = 3, VGEN_P2TR = 4, VGEN_ETHEREUM = 5
};

typedef struct vgen_ctx vgen_ctx;

typedef struct vgen_hit {
    uint8_t key_be[32];      /* secret key, big-endian */
    uint8_t digest[32];      /* hash160 / Ethereum address bytes (20) or tweaked x-only key (32) */
    uint32_t digest_len;
    uint32_t device;         /* lane (index into the device list given at create) that found it */
} vgen_hit;

/* progress callback: cumulative number of valid keys tested so far; called on the calling thread
 * (ProgressCallback, src/scanner.rs:71; src/gpu.rs:1107-1109) */
typedef void (*vgen_progress_fn)(uint64_t keys_scanned, void *user);

int vgen_cuda_device_count(void);
int vgen_cuda_device_name(int device, char *buf, size_t len);

/* devices == NULL or n_devices == 0 selects device 0. keys_per_launch_hint mirrors ScanConfig.gpu_batch_size
 * (src/scanner.rs:40, default 524,288 in the reference); 0 = two resident CTAs per SM (~77.6 M keys / launch). */
int vgen_cuda_create(const int *devices, int n_devices, uint32_t keys_per_launch_hint, vgen_ctx **out);
void vgen_cuda_destroy(vgen_ctx *ctx);
/* message of the last error on this context (ctx == NULL: last error of a failed create on this thread) */
const char *vgen_cuda_last_error(const vgen_ctx *ctx);

/* Pattern::new(pattern, case_insensitive) + lowering for `format`. */
int vgen_cuda_set_pattern(vgen_ctx *ctx, int format, const char *pattern, int case_insensitive);
/* how the current pattern was lowered: "intervals" | "exact" | "dfa" | "all" | "none" */
const char *vgen_cuda_matcher_kind(const vgen_ctx *ctx);

/* This context handles global lanes [shard_index*n_devices, (shard_index+1)*n_devices) of shard_count*n_devices:
 * lets one process per GPU (torchrun) split a range with disjoint start scalars and a common stride. */
int vgen_cuda_set_shard(vgen_ctx *ctx, uint32_t shard_index, uint32_t shard_count);
/* run lane `lane` on a caller-owned cudaStream_t (passed as void*), e.g. torch's current stream */
int vgen_cuda_set_stream(vgen_ctx *ctx, int lane, void *cuda_stream);

/* Scan [start, end] inclusive (end == NULL: up to n-1) with the current pattern. Scalars 0 and >= n are
 * skipped and not counted (src/scanner.rs:294-295). Hits are found in ascending key order, at most
 * max_hits of them (0: unlimited); the first min(found, hits_cap) are written and *n_hits is that number (the rest
 * stay readable through vgen_cuda_scan_fetch_hits).
 * `stop` (may be NULL) is polled after every window; progress (may be NULL) is called after every window. */
int vgen_cuda_scan(vgen_ctx *ctx, const uint8_t start_be[32], const uint8_t *end_be, uint64_t max_hits,
                   const volatile int32_t *stop, vgen_progress_fn progress, void *user,
                   vgen_hit *hits_out, uint64_t hits_cap, uint64_t *n_hits, uint64_t *keys_scanned);

/* The last scan keeps its full confirmed hit list (at most max_hits) inside the context: total count, and a reader for
 * the hits that did not fit the caller's buffer (returns the number written). */
uint64_t vgen_cuda_scan_total_hits(const vgen_ctx *ctx);
uint64_t vgen_cuda_scan_fetch_hits(const vgen_ctx *ctx, uint64_t offset, vgen_hit *out, uint64_t cap);

/* Bulk form of what scan_gpu_with_runner builds per match (src/gpu.rs:1062-1090: address string, WIF, key): for hits
 * [offset, offset+cap) of the last scan writes 32-byte big-endian keys and NUL-terminated address / WIF strings at
 * `stride` (>= 96) bytes per hit. Any of the three outputs may be NULL. Returns the number of hits written. */
uint64_t vgen_cuda_scan_fetch_records(const vgen_ctx *ctx, uint64_t offset, uint64_t cap, uint8_t *keys32, char *addresses, char *wifs, size_t stride);

/* Same records as text, "<key hex> <address> <wif>\n" per hit, formatted on all host threads; at most max_hits hits and
 * what fits in cap bytes. Returns the number of bytes written (NUL-terminated), *n_written = hits written. */
int64_t vgen_cuda_scan_fetch_text(const vgen_ctx *ctx, uint64_t offset, uint64_t max_hits, char *buf, size_t cap, uint64_t *n_written);

/* The same records without a copy: when every hit of the last scan (at most max_hits) has its text record — the scan's confirm
 * workers build them for all but very dense windows — *text points at the library's own buffer holding all of them back to back
 * (NUL-terminated, *len bytes) and 1 is returned. The buffer stays valid until the next scan or set_pattern on this context.
 * Returns 0 (and leaves the outputs untouched) when some record would have to be formatted on demand: use vgen_cuda_scan_fetch_text. */
int vgen_cuda_scan_text(const vgen_ctx *ctx, const char **text, uint64_t *len);

/* generate_batch (src/gpu.rs:660): hash160 of the compressed public key of n sequential scalars, 20 bytes
 * each; zeros for invalid scalars. */
int vgen_cuda_hash160_batch(vgen_ctx *ctx, const uint8_t start_be[32], uint64_t n, uint8_t *out20);
/* same for any format's digest, `stride` bytes per key (20; 32 for P2TR, whose digest is the tweaked witness program) */
int vgen_cuda_digest_batch(vgen_ctx *ctx, int format, const uint8_t start_be[32], uint64_t n, uint8_t *out, uint32_t stride);
/* format value accepted by vgen_cuda_digest_batch only: the P2TR internal key (x-only, even Y) BEFORE the TapTweak, 32 bytes per
 * key — what the reference's dispatch_p2tr / await_result_p2tr return (src/gpu.rs:667-807) */
#define VGEN_P2TR_INTERNAL_KEY 6

/* hash160 batch WITHOUT any allocation. No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same result as vgen_cuda_hash160_batch (hash160 of the compressed public key of n sequential scalars from start_be, 20 bytes
 * each, zeros for scalars 0 and >= the group order) and the same kernels, but it needs no context and everything the kernels
 * work in lives in one file-scope __device__ array of the library (vgen_b200/csrc/static_batch.cu: 17.45 MB, present on every
 * GPU that loads the library until the process exits) instead of cudaMalloc'ed buffers. The call never calls cudaMalloc,
 * cudaFree, cudaHostAlloc or an allocating library routine.
 *   device      CUDA device ordinal (made current by the call)
 *   n           at most VGEN_STATIC_MAX_KEYS; more is rejected with VGEN_ERR_INVALID_ARG and a message, nothing is launched
 *   out20_dev   caller-owned DEVICE memory, n * 20 bytes
 *   cuda_stream cudaStream_t the work is queued on (NULL: the default stream). The call returns once the work is queued; it
 *               synchronises the stream only on the first call per device (table set-up), which therefore must not be made
 *               under graph capture
 *   err         message buffer (may be NULL)
 * All calls on one device share the array, which keeps the last call's contents (the call clears what must start at zero):
 * calls that overlap on the GPU — different streams without an event between them — race. Serialise them. */
#define VGEN_STATIC_MAX_KEYS (1u << 20)
int vgen_cuda_hash160_batch_static(int device, const uint8_t start_be[32], uint32_t n, uint8_t *out20_dev, void *cuda_stream,
                                   char *err, size_t err_cap);

/* Device-resident timing for bench.py: centres set up untimed, then n_windows full-size windows of the current
 * pattern timed with CUDA events on lane 0's stream. */
int vgen_cuda_bench_resident(vgen_ctx *ctx, const uint8_t start_be[32], uint32_t n_windows, float *ms, uint64_t *keys, uint64_t *launches);
/* kernel launches issued by this context so far; keys per full-size window */
uint64_t vgen_cuda_launch_count(const vgen_ctx *ctx);
uint64_t vgen_cuda_window_keys(const vgen_ctx *ctx);
/* lanes of the context: every GPU is driven through 2 streams (each with its own centres, scratch and hit buffers) whose
 * windows overlap, so that a window's late warps do not leave the SMs under-occupied; lanes = 2 x GPUs */
int vgen_cuda_lane_count(const vgen_ctx *ctx);
/* geometry of the P2TR fixed-base table this context has built: 0 none yet; a plain window width in bits (only when VGEN_TAP_BITS
 * forces one, or as the 26-bit fallback); or 0x40000000 | n_hi << 8 | W for signed digits (n_hi windows of W + 1 bits, the rest of
 * W bits): 0x40000111 (15 windows, 67 MB) for short jobs, 0x40000619 (10 windows, 17 GB) for jobs of at least 2^30 keys, 0x4000041c
 * (9 windows, 111.7 GB) from 2^36 keys when the device has the room (VGEN_TAP_BITS: a width, or sW.n_hi); a bigger table is kept
 * once built */
int vgen_cuda_tap_bits(const vgen_ctx *ctx);
/* Counters of the context, for tests and run reports: out[0] devices, out[1] devices that have dropped out after a CUDA error
 * (their windows were taken over by the others; VGEN_FAULT_INJECT=<device index>:<n> makes the n-th window issue of a device
 * fail that way), out[2] windows taken over from a failed device, out[3] windows whose hits exceeded the per-slot hit buffer and
 * were re-run in pieces, out[4] hit records per slot buffer, out[5] confirm worker threads, out[6] guarded device buffers and out[7] how many of their
 * canaries are damaged (both 0 unless the context was created under VGEN_GUARD=1: a 256-byte canary behind every device buffer). */
void vgen_cuda_stats(const vgen_ctx *ctx, uint64_t out[8]);
/* bytes the scans of this context copied host->device (start scalars, launch parameters) and device->host (counters, hit records) */
void vgen_cuda_io_bytes(const vgen_ctx *ctx, uint64_t *h2d, uint64_t *d2h);

/* Host-side string work the reference does per key and this library does per hit. Return the string length. */
int vgen_format_address(int format, const uint8_t *digest, char *out, size_t cap);
int vgen_format_wif(int format, const uint8_t key_be[32], char *out, size_t cap);
/* Pattern::new error check and Pattern::matches on an arbitrary string (1 / 0, <0 on a bad pattern) */
int vgen_pattern_check(const char *pattern, int case_insensitive, char *err, size_t cap);
int vgen_pattern_is_match(const char *pattern, int case_insensitive, const char *text);
/* How (format, pattern) is lowered for the device, without touching a GPU. Writes
 * "<kind>[;<lo16hex>-<hi16hex>]*[;exact=<hex>]" (intervals over the leading 64 digest bits, inclusive);
 * returns the string length. kind as vgen_cuda_matcher_kind. */
/* The window schedule of a scan, without touching a GPU (host arithmetic only; what vgen_cuda_scan follows): the windows
 * that global lane `lane` of `global_lanes` (= shard_count x GPUs per context) takes from [start, end] on devices with
 * `sm_count` SMs. Writes up to cap windows (first key, big-endian, 32 bytes each; keys inside the range) and returns the
 * total number of windows of that lane. can_end_early != 0 selects the ramped schedule used for hit-limited dense scans.
 * Lanes own disjoint windows whose union is exactly the valid part of the range. */
int64_t vgen_plan_windows(const uint8_t start_be[32], const uint8_t end_be[32], int sm_count, uint32_t keys_per_launch_hint,
                          uint32_t global_lanes, uint32_t lane, int can_end_early, uint8_t *first_keys_be, uint64_t *n_keys, uint64_t cap);

/* The dense DFA the device runs for (format, pattern): table[n_states*n_classes] (row-major next state), accept[n_states],
 * sym_class[64] (encoder alphabet index -> class), info = {n_states, n_classes, start state after the fixed address head,
 * flags}. Returns VGEN_ERR_BUFFER (info still filled) when a buffer is too small. For inspection and tests. */
int vgen_pattern_dfa(int format, const char *pattern, int case_insensitive, uint16_t *table, size_t table_cap,
                     uint8_t *accept, size_t accept_cap, uint8_t sym_class[64], uint32_t info[4]);
int vgen_pattern_lowering(int format, const char *pattern, int case_insensitive, char *out, size_t cap);


/* ---- CLI support (host only): the reference's pattern lint and difficulty heuristic
 * (Pattern::validate_charset, src/pattern.rs:48-175; Pattern::estimate_difficulty, src/pattern.rs:177-253), its provider
 * lookup (provider::resolve / build_pattern / build_exact_pattern, src/provider.rs:12-63; only an embedded subset of
 * boha's b1000 collection) and the key parser of `verify` (src/lib.rs:387-401). */
int vgen_pattern_validate_charset(int format, const char *pattern, int case_insensitive, char *invalid_chars_out, size_t cap);
uint64_t vgen_pattern_estimate_difficulty(int format, const char *pattern, int case_insensitive);
/* 1 = resolved (address, format, key range filled), 0 = not a provider reference, <0 = provider error (message in err) */
int vgen_provider_resolve(const char *ref, char *address, size_t cap, int *format, uint8_t start_be[32], uint8_t end_be[32],
                          int *has_range, char *err, size_t err_cap);
/* "^<escaped prefix>" for prefix_len > 0, "^<escaped address>$" for prefix_len <= 0 */
int vgen_provider_pattern(const char *address, int prefix_len, char *out, size_t cap);
/* WIF or hex -> 32-byte key; VGEN_ERR_INVALID_ARG if neither, VGEN_ERR_RANGE if the scalar is 0 or >= n */
int vgen_parse_key(const char *text, uint8_t key_be[32], int *is_wif, int *compressed);

#ifdef __cplusplus
}
#endif
#endif
