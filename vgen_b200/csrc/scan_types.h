// Plain-C types shared by the device code (scan_core.cuh) and the host engine (engine.cpp).
#pragma once
#include <stdint.h>

enum { VG_FMT_P2PKH = 0, VG_FMT_P2PKH_UNCOMPRESSED = 1, VG_FMT_P2WPKH = 2, VG_FMT_P2SH_P2WPKH = 3, VG_FMT_P2TR = 4, VG_FMT_ETHEREUM = 5 };
enum { VG_MATCH_ALL = 0, VG_MATCH_INTERVALS = 1, VG_MATCH_EXACT = 2, VG_MATCH_DFA = 3 };
enum { VG_MODE_MATCH = 0, VG_MODE_DUMP = 1 };

#define VG_MAX_INTERVALS 32
#define VG_DIGEST_WORDS 8

// ---- geometry of the P2TR fixed-base table (taproot.cuh), shared by the kernels and the host engine ----
// The geometry word `g` (vg_scan_params::tap_bits):
//   plain  (VG_TAP_SIGNED clear): uniform unsigned windows of g bits; window j, digit v at entry (j << g) + v holds v * 2^(g*j) * G.
//   signed (VG_TAP_SIGNED set)  : W = g & 0xff, n_hi = (g >> 8) & 0xff: the first n_hi windows are W + 1 bits wide, the others W, and
//          their widths add up to exactly 256. Digits are signed, |d| <= 2^(width-1); a window holds the 2^(width-1) + 1 multiples
//          0 .. 2^(width-1) of its base point and a negative digit takes the entry with y negated: half the entries per window,
//          which is what lets 9 windows (4 x 29 + 5 x 28 bits) fit: 1.745 G entries = 111.7 GB of the 180 GB.
#if defined(__CUDACC__)
#define VG_HD __host__ __device__ __forceinline__
#else
#define VG_HD static inline
#endif
#define VG_TAP_SIGNED 0x40000000u
#define VG_TAP_GEOM_SIGNED(W, n_hi) (VG_TAP_SIGNED | ((uint32_t)(n_hi) << 8) | (uint32_t)(W))
VG_HD uint32_t vg_tap_windows(uint32_t g) {
    if (g & VG_TAP_SIGNED) return (256u - ((g >> 8) & 0xffu)) / (g & 0xffu);
    return (256u + g - 1u) / g;
}
// window j (j == number of windows: one past the end): its first bit, its width and the table index of its entry 0
VG_HD void vg_tap_window(uint32_t g, uint32_t j, uint32_t *off, uint32_t *width, unsigned long long *first) {
    if (g & VG_TAP_SIGNED) {
        const uint32_t W = g & 0xffu, nh = (g >> 8) & 0xffu, jh = j < nh ? j : nh;
        *width = W + (j < nh ? 1u : 0u);
        *off = j * W + jh;
        *first = (unsigned long long)jh * ((1ull << W) + 1ull) + (unsigned long long)(j - jh) * ((1ull << (W - 1u)) + 1ull);
    } else { *width = g; *off = j * g; *first = (unsigned long long)j << g; }
}
VG_HD unsigned long long vg_tap_entries(uint32_t g) {
    uint32_t off, width; unsigned long long first;
    vg_tap_window(g, vg_tap_windows(g), &off, &width, &first);
    return first;
}
// is g a geometry the kernels can run?
VG_HD int vg_tap_geom_valid(uint32_t g) {
    if (g & VG_TAP_SIGNED) {
        const uint32_t W = g & 0xffu, nh = (g >> 8) & 0xffu;
        return (g & ~(VG_TAP_SIGNED | 0xffffu)) == 0 && W >= 3u && W <= 30u && nh < 256u && (256u - nh) % W == 0u && nh <= (256u - nh) / W;
    }
    return g >= 4u && g <= 26u;
}

struct vg_hit_rec {          // 40 bytes
    uint32_t idx_lo, idx_hi; // key index inside the window
    uint32_t digest[VG_DIGEST_WORDS];   // digest bytes as little-endian words (byte 4i+j = byte j of digest[i])
};
struct vg_dfa_dev {
    const uint16_t *table;   // [n_states][n_classes] next state
    const uint8_t *accept;   // [n_states] bit0: accepting if input ends here
    uint32_t n_states, n_classes;
    uint32_t start;          // start state, already advanced over the fixed head of the address ("bc1q", "0x")
    uint32_t flags;          // bit 0: the DFA tells a-f from A-F (Ethereum: EIP-55 case must be computed on the device)
    uint8_t sym_class[64];   // alphabet index of the format's encoder (base58 digit / bech32 value / hex nibble, +16 = upper case) -> class
    uint32_t table_u16, accept_bytes;   // sizes of the two arrays as uploaded / staged (the folded DFA below is appended to them)
    // Ethereum, case-sensitive patterns only: the same pattern compiled case-insensitively, appended behind the exact DFA.
    // It runs on the lower-case hex first; only keys it accepts pay the EIP-55 checksum Keccak (match.cuh vg_match_eth).
    uint32_t fold_table, fold_accept;   // offsets of its table (uint16 units) / accept bytes; fold_table == 0: none
    uint32_t fold_classes, fold_start;
    uint8_t fold_sym[16];               // nibble -> class of the folded DFA
    // Segwit formats, patterns that are a literal suffix of at most 6 characters (dead$): the match lies inside the 6 checksum
    // symbols, so (checksum & chk_mask) == chk_value is a necessary condition that needs no DFA step (match.cuh
    // vg_match_segwit); chk_mask == 0: none. Ethereum, literal hex suffixes of at most 8 digits (F00$): the same pair over the last
    // 32 digest bits (big-endian), tested before the DFAs and the EIP-55 hash (vg_match_eth).
    uint32_t chk_mask, chk_value;
    // Base58 formats, patterns that are a literal suffix of at most 5 characters (Cat$): the last k characters of the address are
    // the 25-byte Base58Check payload mod 58^k, so (payload mod b58_mod) in b58_vals[0..b58_nvals) is a necessary condition that
    // needs neither the 28 divisions of the radix conversion nor a DFA step (match.cuh vg_match_base58check); b58_mod == 0: none.
    // b58_w[j] = 2^(16 j) mod b58_mod for the 13 half-words of the payload, b58_mu = floor(2^64 / b58_mod) (Barrett).
    uint32_t b58_mod, b58_nvals;
    uint32_t b58_vals[8];
    uint32_t b58_w[13];
    uint32_t b58_mu_lo, b58_mu_hi;
};
struct vg_matcher_dev {
    int kind;
    int n_intervals;
    uint32_t lo_hi[VG_MAX_INTERVALS][4];   // {lo_hi32, lo_lo32, hi_hi32, hi_lo32}: inclusive range of the first 8 digest bytes (big-endian)
    uint32_t exact[VG_DIGEST_WORDS];       // little-endian words, as produced by the hashes
    int exact_words;
    int confirm_dfa;                       // run the DFA on candidates that pass the prefilter
    vg_dfa_dev dfa;
};
struct vg_scan_params {
    uint32_t n_threads, half;
    uint32_t *centers;        // [16][n_threads]: x limbs 0..7 then y limbs 0..7 (canonical); all-zero = no valid centre
    uint32_t *scratch;        // [(half+1)][8][n_threads] prefix products
    const uint32_t *gtab;     // [half][16]: (i+1)*G affine canonical, x limbs then y limbs
    uint32_t stride_pt[16];   // W*G affine canonical
    unsigned long long limit; // window keys with index < limit are inside the scan range
    vg_hit_rec *hits; uint32_t *hit_count; uint32_t hit_cap;
    uint32_t *degen_list; uint32_t *degen_count; uint32_t degen_cap;
    uint32_t dfa_smem_bytes;  // shared-memory bytes for the DFA table + accept flags (0: read them from global memory)
    uint8_t *dump; uint32_t dump_stride;   // VG_MODE_DUMP: digest bytes per key index
    // P2TR only (taproot.cuh): stage 1 writes the even-Y internal key of window key i to points[16*i .. 16*i+15]
    // (x limbs, y limbs; all-zero = invalid), stage 2 walks tap_table (16 words per entry; geometry word tap_bits, see above)
    uint32_t *points; const uint32_t *tap_table; uint32_t tap_bits;
    uint32_t tap_base[8];     // scalar of window key 0 (little-endian limbs): stage 2 derives Q = (t +/- k) * G from the key's scalar
    // checked builds (-DVG_CHECKED, tools/check_build.sh): [0] = code of the first failed device-side bounds check, [1] = how many
    // failed. compute-sanitizer is closed on this pool, so the index arithmetic of every kernel carries its own asserts instead.
    uint32_t *violations;
    vg_matcher_dev m;
};

