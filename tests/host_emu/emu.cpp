// TEST HARNESS: compiles the product's device headers for the host (VG_HOST_EMU) so the field, curve,
// hash and walk logic can be checked against the oracle on a machine without a GPU.
// Not part of the product; never used on the GPU path.
#define VG_HOST_EMU 1
#define VG_B58_TABLE 1     // the table-driven Base58 conversion is off in the product build; keep it checked here
#include <vector>
#include <cstring>
#include "../../vgen_b200/csrc/taproot.cuh"

static const uint32_t G_XY[16] = {
    0x16F81798, 0x59F2815B, 0x2DCE28D9, 0x029BFCDB, 0xCE870B07, 0x55A06295, 0xF9DCBBAC, 0x79BE667E,
    0xFB10D4B8, 0x9C47D08F, 0xA6855419, 0xFD17B448, 0x0E1108A8, 0x5DA4FBFC, 0x26A3C465, 0x483ADA77};
static std::vector<uint32_t> g_pow2;
static void ensure_pow2() { if (g_pow2.empty()) { g_pow2.resize(256 * 16); std::vector<uint32_t> scratch(256 * 32); vg_build_pow2(g_pow2.data(), G_XY, scratch.data()); } }

extern "C" {
void emu_fe_mul(const uint32_t *a, const uint32_t *b, uint32_t *r) { fe x, y, z; memcpy(x.v, a, 32); memcpy(y.v, b, 32); fe_mul(z, x, y); fe_canon(z); memcpy(r, z.v, 32); }
void emu_fe_sqr_fast(const uint32_t *a, uint32_t *r) { fe x, z; memcpy(x.v, a, 32); fe_sqr_fast(z, x); fe_canon(z); memcpy(r, z.v, 32); }
void emu_fe_add(const uint32_t *a, const uint32_t *b, uint32_t *r) { fe x, y, z; memcpy(x.v, a, 32); memcpy(y.v, b, 32); fe_add(z, x, y); fe_canon(z); memcpy(r, z.v, 32); }
void emu_fe_sub(const uint32_t *a, const uint32_t *b, uint32_t *r) { fe x, y, z; memcpy(x.v, a, 32); memcpy(y.v, b, 32); fe_sub(z, x, y); fe_canon(z); memcpy(r, z.v, 32); }
void emu_fe_inv(const uint32_t *a, uint32_t *r) { fe x, z; memcpy(x.v, a, 32); fe_inv(z, x); fe_canon(z); memcpy(r, z.v, 32); }
int emu_scalar_mul(const uint32_t *k, uint32_t *x, uint32_t *y) {
    ensure_pow2(); fe fx, fy; bool ok = scalar_is_valid(k) && scalar_mul_pow2(fx, fy, k, g_pow2.data());
    memcpy(x, fx.v, 32); memcpy(y, fy.v, 32); return ok ? 0 : -1;
}
// the device's 8-bit windowed k*G (ec.cuh scalar_mul_G with w8 = true); builds the win8 part of the table on first use
static std::vector<uint32_t> g_gtable;
int emu_scalar_mul_w8(const uint32_t *k, uint32_t *x, uint32_t *y) {
    ensure_pow2();
    if (g_gtable.empty()) {
        g_gtable.assign(VG_GTABLE_WORDS, 0);
        memcpy(g_gtable.data(), g_pow2.data(), VG_GTABLE_POW2_WORDS * 4);
        for (uint32_t i = 0; i < 32 * 256; i++) {
            uint32_t j = i >> 8, v = i & 255u, kk[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            if (!v) continue;
            kk[j >> 2] = v << (8 * (j & 3));
            fe px, py; scalar_mul_pow2(px, py, kk, g_pow2.data());
            memcpy(&g_gtable[VG_GTABLE_POW2_WORDS + (size_t)i * 16], px.v, 32); memcpy(&g_gtable[VG_GTABLE_POW2_WORDS + (size_t)i * 16 + 8], py.v, 32);
        }
    }
    fe fx, fy; bool ok = scalar_is_valid(k) && scalar_mul_G(fx, fy, k, g_gtable.data(), true);
    memcpy(x, fx.v, 32); memcpy(y, fy.v, 32); return ok ? 0 : -1;
}
void emu_digest(int fmt, const uint32_t *x, const uint32_t *y, uint32_t *dg) {
    fe fx, fy; memcpy(fx.v, x, 32); memcpy(fy.v, y, 32);
    switch (fmt) {
        case 0: vg_digest_of_point<0>(dg, fx, fy); break; case 1: vg_digest_of_point<1>(dg, fx, fy); break;
        case 2: vg_digest_of_point<2>(dg, fx, fy); break; case 3: vg_digest_of_point<3>(dg, fx, fy); break;
        case 4: vg_digest_of_point<4>(dg, fx, fy); break; default: vg_digest_of_point<5>(dg, fx, fy); break;
    }
}
// on-device encoder + DFA of match.cuh for one digest: returns 1 when the device matcher accepts
int emu_confirm(int fmt, const uint32_t *dg, const uint16_t *table, const uint8_t *accept, const uint8_t *sym_class,
                uint32_t n_classes, uint32_t start, uint32_t flags) {
    vg_matcher_dev m; memset(&m, 0, sizeof m);
    m.dfa.n_classes = n_classes; m.dfa.start = start; m.dfa.flags = flags;
    vg_dfa_view dv; dv.table = table; dv.accept = accept; dv.sym = sym_class;
    if (fmt == 2 || fmt == 4) {
        // segwit formats: the stepwise polymod and the table-driven checksum (what k_walk / k_taptweak use) must agree
        static uint32_t tab32[32 * 32 + 1], tab52[52 * 32 + 1]; static bool built = false;
        if (!built) { vg_bech32_fill_table<32>(tab32, 0u, VG_BECH32_CONST, 0, 1); vg_bech32_fill_table<52>(tab52, 1u, VG_BECH32M_CONST, 0, 1); built = true; }
        int a = fmt == 2 ? vg_confirm<2>(m, dv, dg) : vg_confirm<4>(m, dv, dg);
        dv.poly = fmt == 2 ? tab32 : tab52;
        int b = fmt == 2 ? vg_confirm<2>(m, dv, dg) : vg_confirm<4>(m, dv, dg);
        return a == b ? a : -1;
    }
    if (fmt == 0 || fmt == 1 || fmt == 3) {
        // Base58 formats: long division and the table-driven radix conversion must agree
        static uint32_t b58[VG_B58_TABLE_WORDS]; static bool built58 = false;
        if (!built58) { vg_b58_fill_table(b58, 0, 1); built58 = true; }
        int a = fmt == 3 ? vg_confirm<3>(m, dv, dg) : vg_confirm<0>(m, dv, dg);
        dv.b58 = b58;
        int b = fmt == 3 ? vg_confirm<3>(m, dv, dg) : vg_confirm<0>(m, dv, dg);
        return a == b ? a : -1;
    }
    switch (fmt) {
        case 0: return vg_confirm<0>(m, dv, dg); case 1: return vg_confirm<1>(m, dv, dg); case 2: return vg_confirm<2>(m, dv, dg);
        case 3: return vg_confirm<3>(m, dv, dg); case 4: return vg_confirm<4>(m, dv, dg); default: return vg_confirm<5>(m, dv, dg);
    }
}
// segwit matcher with the checksum-suffix prefilter (vg_dfa_dev::chk_mask / chk_value) on the table path
int emu_confirm_chk(int fmt, const uint32_t *dg, const uint16_t *table, const uint8_t *accept, const uint8_t *sym_class,
                    uint32_t n_classes, uint32_t start, uint32_t flags, uint32_t chk_mask, uint32_t chk_value) {
    vg_matcher_dev m; memset(&m, 0, sizeof m);
    m.dfa.n_classes = n_classes; m.dfa.start = start; m.dfa.flags = flags; m.dfa.chk_mask = chk_mask; m.dfa.chk_value = chk_value;
    vg_dfa_view dv; dv.table = table; dv.accept = accept; dv.sym = sym_class;
    static uint32_t tab32[32 * 32 + 1], tab52[52 * 32 + 1]; static bool built = false;
    if (!built) { vg_bech32_fill_table<32>(tab32, 0u, VG_BECH32_CONST, 0, 1); vg_bech32_fill_table<52>(tab52, 1u, VG_BECH32M_CONST, 0, 1); built = true; }
    if (fmt == 5) return vg_confirm<5>(m, dv, dg);       // Ethereum: mask over the last 32 digest bits
    dv.poly = fmt == 2 ? tab32 : tab52;
    return fmt == 2 ? vg_confirm<2>(m, dv, dg) : vg_confirm<4>(m, dv, dg);
}
// Base58 matcher with the literal-suffix residue prefilter (vg_dfa_dev::b58_mod ...): 13 half-word weights, Barrett constant
int emu_confirm_b58res(int fmt, const uint32_t *dg, const uint16_t *table, const uint8_t *accept, const uint8_t *sym_class,
                       uint32_t n_classes, uint32_t start, uint32_t flags, uint32_t mod, uint32_t nvals, const uint32_t *vals,
                       const uint32_t *w, uint32_t mu_lo, uint32_t mu_hi) {
    vg_matcher_dev m; memset(&m, 0, sizeof m);
    m.dfa.n_classes = n_classes; m.dfa.start = start; m.dfa.flags = flags;
    m.dfa.b58_mod = mod; m.dfa.b58_nvals = nvals; m.dfa.b58_mu_lo = mu_lo; m.dfa.b58_mu_hi = mu_hi;
    memcpy(m.dfa.b58_vals, vals, 4 * (nvals < 8 ? nvals : 8)); memcpy(m.dfa.b58_w, w, 4 * 13);
    vg_dfa_view dv; dv.table = table; dv.accept = accept; dv.sym = sym_class;
    return fmt == 3 ? vg_confirm<3>(m, dv, dg) : vg_confirm<0>(m, dv, dg);
}
// Ethereum matcher with the case-folded DFA appended behind the exact one (lowering.cpp): table / accept hold both
int emu_confirm_eth_fold(const uint32_t *dg, const uint16_t *table, const uint8_t *accept, const uint8_t *sym_class,
                         uint32_t n_classes, uint32_t start, uint32_t flags, uint32_t fold_table, uint32_t fold_accept,
                         uint32_t fold_classes, uint32_t fold_start, const uint8_t *fold_sym) {
    vg_matcher_dev m; memset(&m, 0, sizeof m);
    m.dfa.n_classes = n_classes; m.dfa.start = start; m.dfa.flags = flags;
    m.dfa.fold_table = fold_table; m.dfa.fold_accept = fold_accept; m.dfa.fold_classes = fold_classes; m.dfa.fold_start = fold_start;
    memcpy(m.dfa.fold_sym, fold_sym, 16);
    vg_dfa_view dv; dv.table = table; dv.accept = accept; dv.sym = sym_class;
    return vg_confirm<5>(m, dv, dg);
}
// P2TR stage 2 (taproot.cuh) over n keys (n a multiple of VG_TAP_GROUP): key i has the scalar base8 + i and its public key (x, y) in
// points[16*i ..] (all-zero = invalid), as stage 1 leaves them. Writes the 32-byte witness programs. bits = window width of the
// fixed-base table (4 keeps the host build cheap).
static std::vector<uint32_t> g_tap; static uint32_t g_tap_bits = 0;
int emu_tap_scalar_form() { return VG_TAP_SCALAR; }
// table geometry (scan_types.h): out = {valid, windows, entries, first bit past the top window}
void emu_tap_geom(uint32_t g, unsigned long long *out) {
    out[0] = (unsigned long long)vg_tap_geom_valid(g);
    if (!out[0]) return;
    out[1] = vg_tap_windows(g); out[2] = vg_tap_entries(g);
    uint32_t off, width; unsigned long long first;
    vg_tap_window(g, (uint32_t)out[1] - 1, &off, &width, &first);
    out[3] = off + width;
}
void emu_scalar_add_mod_n(const uint32_t *a, const uint32_t *b, uint32_t *r) { vg_scalar_add_mod_n(r, a, b); }
void emu_scalar_neg_n(const uint32_t *k, uint32_t *r) { vg_scalar_neg_n(r, k); }
void emu_tap_stage2(uint32_t *points, const uint32_t *base8, uint32_t n, uint32_t bits, uint8_t *out32) {
    ensure_pow2();
    if (g_tap_bits != bits) {          // bits: a geometry word (scan_types.h): plain 4..26, or VG_TAP_SIGNED | n_hi << 8 | W
        g_tap.assign((size_t)vg_tap_entries(bits) * 16, 0);
        for (uint32_t j = 0; j < vg_tap_windows(bits); j++) {
            uint32_t off, width, o2, w2; unsigned long long first, next;
            vg_tap_window(bits, j, &off, &width, &first);
            vg_tap_window(bits, j + 1, &o2, &w2, &next);
            for (unsigned long long i = first; i < next; i++) {
                uint32_t k[8];
                if (!vg_tap_digit_scalar(k, off, (uint32_t)(i - first))) continue;
                fe x, y; scalar_mul_pow2(x, y, k, g_pow2.data());
                memcpy(&g_tap[i * 16], x.v, 32); memcpy(&g_tap[i * 16 + 8], y.v, 32);
            }
        }
        g_tap_bits = bits;
    }
    vg_scan_params P; memset(&P, 0, sizeof P);
    uint32_t hitc = 0;
    P.points = points; P.tap_table = g_tap.data(); P.tap_bits = bits; P.limit = n; P.dump = out32; P.dump_stride = 32; P.hit_count = &hitc;
    memcpy(P.tap_base, base8, 32);
    vg_dfa_view dv; dv.table = nullptr; dv.accept = nullptr; dv.sym = nullptr;
    memset(out32, 0, (size_t)n * 32);
    for (uint32_t g = 0; g < n / VG_TAP_GROUP; g++) vg_tap_thread<VG_MODE_DUMP>(P, dv, (unsigned long long)g * VG_TAP_GROUP);
}
void emu_keccak_hex40(const uint32_t *m, uint32_t *out) { keccak_hex40(out, m); }

// Walk n_windows windows of T threads x 2H keys from scalar `start`, dumping `stride` digest bytes per key.
// Returns the number of degenerate thread-batches seen.
int emu_walk_dump(int fmt, const uint32_t *start, uint32_t T, uint32_t H, uint32_t n_windows, uint32_t stride, uint8_t *out) {
    ensure_pow2();
    const uint64_t B = 2ull * H, W = (uint64_t)T * B;
    std::vector<uint32_t> gtab(H * 16), centers(16 * T), scratch((size_t)(H + 1) * 8 * T), colbuf(16);
    for (uint32_t i = 0; i < H; i++) {
        uint32_t k[8] = {i + 1, 0, 0, 0, 0, 0, 0, 0};
        vg_point_of_scalar(colbuf.data(), 1, 0, k, g_pow2.data());
        memcpy(&gtab[i * 16], colbuf.data(), 64);
    }
    vg_scan_params P; memset(&P, 0, sizeof P);
    uint32_t wk[8] = {(uint32_t)W, (uint32_t)(W >> 32), 0, 0, 0, 0, 0, 0};
    vg_point_of_scalar(colbuf.data(), 1, 0, wk, g_pow2.data());
    memcpy(P.stride_pt, colbuf.data(), 64);
    for (uint32_t t = 0; t < T; t++) {
        uint32_t k[8]; scalar_add_u64(k, start, (uint64_t)t * B + H);
        vg_point_of_scalar(centers.data(), T, t, k, g_pow2.data());
    }
    uint32_t hitc = 0, degc = 0; std::vector<uint32_t> degl(T);
    P.n_threads = T; P.half = H; P.centers = centers.data(); P.scratch = scratch.data(); P.gtab = gtab.data();
    P.limit = W; P.hit_count = &hitc; P.degen_count = &degc; P.degen_list = degl.data(); P.degen_cap = T;
    P.dump_stride = stride;
    std::vector<uint32_t> points(fmt == 4 ? (size_t)W * 16 : 0);
    P.points = points.data();
    vg_dfa_view dv; dv.table = nullptr; dv.accept = nullptr; dv.sym = nullptr;
    for (uint32_t w = 0; w < n_windows; w++) {
        P.dump = out + (size_t)w * W * stride;
        for (uint32_t t = 0; t < T; t++) {
            switch (fmt) {
                case 0: vg_walk_thread<0, VG_MODE_DUMP>(P, dv, gtab.data(), t); break;
                case 1: vg_walk_thread<1, VG_MODE_DUMP>(P, dv, gtab.data(), t); break;
                case 3: vg_walk_thread<3, VG_MODE_DUMP>(P, dv, gtab.data(), t); break;
                case 4: vg_walk_thread<4, VG_MODE_DUMP>(P, dv, gtab.data(), t); break;
                default: vg_walk_thread<5, VG_MODE_DUMP>(P, dv, gtab.data(), t); break;
            }
        }
        if (fmt == 4) {
            uint32_t wb[8]; scalar_add_u64(wb, start, (uint64_t)w * W);
            emu_tap_stage2(points.data(), wb, (uint32_t)W, 4, P.dump);      // stride must be 32
        }
    }
    return (int)degc;
}
}
