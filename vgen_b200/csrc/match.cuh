// Device-side exact matching for general regexes: the address string of each key is produced ON THE GPU
// (Base58Check, Bech32, EIP-55 hex) symbol by symbol and fed straight into a dense DFA whose transition table is
// staged in shared memory. No string is ever materialised: every encoder yields alphabet indices, and the host
// pre-maps alphabet index -> DFA character class (vg_dfa_dev::sym_class), and pre-advances the DFA start state over
// the fixed part of the address ("bc1q", "0x").
//
// Replaces the reference's per-key host work: hash160 -> bitcoin::Address string -> regex (src/gpu.rs:1030-1093),
// and to_checksum_address (src/address.rs:176-198). Semantics = Pattern::matches (src/pattern.rs:43-45): regex
// search over the whole address string. Host-side confirmation of every device hit (engine.cpp) keeps the reported
// match set exact even so; tests check that device and host agree in both directions.
#pragma once
#include "scan_core.cuh"

#ifdef VG_HOST_EMU
#define VG_CLZ(x) ((uint32_t)__builtin_clz(x))
static inline unsigned long long vg_umul64hi(unsigned long long a, unsigned long long b) { return (unsigned long long)(((unsigned __int128)a * b) >> 64); }
#else
#define VG_CLZ(x) ((uint32_t)__clz((int)(x)))
__device__ __forceinline__ unsigned long long vg_umul64hi(unsigned long long a, unsigned long long b) { return __umul64hi(a, b); }
#endif

struct vg_dfa_view {
    const uint16_t *table;     // [n_states][n_classes], shared memory when it fits, else global
    const uint8_t *accept;     // [n_states]
    const uint8_t *sym;        // [64] alphabet index -> class (vg_dfa_dev::sym_class)
    const uint32_t *poly = nullptr;   // Bech32 checksum table in shared memory (vg_bech32_fill_table), or null: stepwise polymod
    const uint32_t *b58 = nullptr;    // radix-conversion table in shared memory (vg_b58_fill_table), or null: long division
    uint32_t n_table = 0, n_accept = 0;   // entries of table / accept (checked builds)
    uint32_t *viol = nullptr;             // vg_scan_params::violations (checked builds)
};

VG_DEV uint32_t vg_dfa_step(const vg_dfa_view &v, uint32_t n_classes, uint32_t st, uint32_t cls) {
    VG_ASSERT_AT(v.viol, st * n_classes + cls < v.n_table && cls < n_classes, 7);
    return v.table[st * n_classes + cls];
}
VG_DEV bool vg_dfa_accepts(const vg_dfa_view &v, uint32_t st) {
    VG_ASSERT_AT(v.viol, st < v.n_accept, 8);
    return v.accept[st] != 0;
}

// ---- SHA-256 of exactly 32 bytes given as 8 big-endian words (second half of the Base58Check checksum) ----
VG_DEV uint32_t sha256_32_first_word(const uint32_t in[8]) {
    uint32_t w[16];
    VG_UNROLL for (int i = 0; i < 8; i++) w[i] = in[i];
    w[8] = 0x80000000u;
    VG_UNROLL for (int i = 9; i < 15; i++) w[i] = 0;
    w[15] = 256;
    uint32_t st[8];
    sha256_iv(st);
    sha256_compress<0xFF00u>(st, w);
    return st[0];
}

// ---- Base58 by table ------------------------------------------------------------------------------------------------
// The 25-byte payload N = sum_i nib_i * 16^i (50 nibbles). tab[(i*9 + g)*16 + v] = g-th base-58^4 digit of v * 16^i, so the nine
// base-58^4 groups of N are column sums of 50 table rows (each sum < 50 * 58^4 < 2^30) followed by one carry pass: shared-memory
// loads and adds in place of the 28 64-bit divisions of the long-division path. Groups that are zero for every v (v * 16^i <
// 58^(4g)) are skipped at compile time: 243 loads per key. A row is 16 consecutive words, so a warp's loads never conflict.
// MEASURED (B200, r01, profiles/b58_table_r01.log): not faster than the long division - p2pkh Cat$ 2,721 against 2,780 Mkeys/s; the
// 243 loads and their address arithmetic cost what the divisions did, and the walk kernels go from 120 to 128 registers. Kept behind
// VG_B58_TABLE (default 0; the host-emulation tests build it with 1 and check it against the division path and the oracle).
#ifndef VG_B58_TABLE
#define VG_B58_TABLE 0
#endif
#define VG_B58_R 11316496u      // 58^4
#define VG_B58_GROUPS 9         // 58^36 > 2^200
#define VG_B58_NIBBLES 50
#define VG_B58_TABLE_WORDS (VG_B58_NIBBLES * VG_B58_GROUPS * 16)
VG_DEV void vg_b58_fill_table(uint32_t *tab, uint32_t tid, uint32_t nthreads) {
    for (uint32_t e = tid; e < VG_B58_NIBBLES * 16u; e += nthreads) {
        const uint32_t i = e >> 4, val = e & 15u;
        uint32_t d[VG_B58_GROUPS];
        d[0] = val;
        VG_UNROLL for (int g = 1; g < VG_B58_GROUPS; g++) d[g] = 0;
        for (uint32_t k = 0; k < i; k++) {            // times 16, i times
            uint32_t carry = 0;
            VG_UNROLL for (int g = 0; g < VG_B58_GROUPS; g++) {
                uint32_t t = d[g] * 16u + carry;      // < 2^28
                carry = t / VG_B58_R;
                d[g] = t - carry * VG_B58_R;
            }
        }
        VG_UNROLL for (int g = 0; g < VG_B58_GROUPS; g++) tab[(i * VG_B58_GROUPS + g) * 16 + val] = d[g];
    }
}

// checksum = first 4 bytes of SHA256(SHA256(version || h160)); B = h160 as big-endian words
template <int VERSION>
VG_DEV uint32_t vg_base58check_checksum(const uint32_t B[5]) {
    uint32_t msg[15];
    msg[0] = ((uint32_t)VERSION << 24) | (B[0] >> 8);
    VG_UNROLL for (int i = 1; i < 5; i++) msg[i] = (B[i - 1] << 24) | (B[i] >> 8);
    msg[5] = B[4] << 24;
    VG_UNROLL for (int i = 6; i < 15; i++) msg[i] = 0;
    uint32_t h1[8];
    sha256_short<21>(h1, msg);
    return sha256_32_first_word(h1);
}
// leading zero BYTES of the payload become leading '1' symbols (alphabet index 0); n = payload limbs, n[6] = version
template <int VERSION>
VG_DEV uint32_t vg_base58_leading_ones(const uint32_t n[7]) {
    uint32_t zeros = 0;
    if (VERSION == 0) {
        zeros = 1;
        bool run = true;
        VG_UNROLL for (int i = 5; i >= 0; i--) {
            if (run) {
                if (n[i] == 0) zeros += 4;
                else { zeros += (VG_CLZ(n[i]) >> 3); run = false; }
            }
        }
    }
    return zeros;
}
template <int VERSION>
VG_DEV bool vg_match_base58check_tab(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t n[7]) {
    uint32_t sum[VG_B58_GROUPS];
    VG_UNROLL for (int g = 0; g < VG_B58_GROUPS; g++) sum[g] = 0;
    VG_UNROLL for (int i = 0; i < VG_B58_NIBBLES; i++) {
        if (i >= 48 && ((VERSION >> (4 * (i - 48))) & 15) == 0) continue;      // a zero nibble adds nothing
        const uint32_t nib = (n[i >> 3] >> (4 * (i & 7))) & 15u;
        VG_UNROLL for (int g = 0; g < VG_B58_GROUPS; g++)
            if (2343 * g <= 100 * (4 * i + 4))                                   // 58^(4g) <= 16^(i+1): log2(58^4) = 23.43
                sum[g] += v.b58[(i * VG_B58_GROUPS + g) * 16 + nib];
    }
    uint32_t carry = 0;
    VG_UNROLL for (int g = 0; g < VG_B58_GROUPS; g++) {
        uint32_t t = sum[g] + carry;
        carry = t / VG_B58_R;
        sum[g] = t - carry * VG_B58_R;
    }
    const uint32_t nc = m.dfa.n_classes;
    uint32_t st = m.dfa.start;
    const uint32_t c_one = v.sym[0];
    const uint32_t zeros = vg_base58_leading_ones<VERSION>(n);
    for (uint32_t z = 0; z < zeros; z++) st = vg_dfa_step(v, nc, st, c_one);
    bool started = false;
    VG_UNROLL for (int r = VG_B58_GROUPS - 1; r >= 0; r--) {
        uint32_t g = sum[r];
        uint32_t d3 = g / 195112u;   g -= d3 * 195112u;       // 58^3
        uint32_t d2 = g / 3364u;     g -= d2 * 3364u;         // 58^2
        uint32_t d1 = g / 58u;       uint32_t d0 = g - d1 * 58u;
        uint32_t ds[4] = {d3, d2, d1, d0};
        VG_UNROLL for (int k = 0; k < 4; k++) {
            started = started || ds[k] != 0;
            if (started) st = vg_dfa_step(v, nc, st, v.sym[ds[k]]);
        }
    }
    return vg_dfa_accepts(v, st);
}

// Base58Check(version || h160): runs the DFA over the address symbols. dg = h160 as little-endian words.
template <int VERSION>
VG_DEV bool vg_match_base58check(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    uint32_t B[5];
    VG_UNROLL for (int i = 0; i < 5; i++) B[i] = vg_bswap(dg[i]);
    uint32_t ck = vg_base58check_checksum<VERSION>(B);
    // 25-byte payload as a 200-bit big-endian number in 7 limbs, n[6] = top 8 bits = version
    uint32_t n[7];
    n[6] = (uint32_t)VERSION;
    n[5] = B[0]; n[4] = B[1]; n[3] = B[2]; n[2] = B[3]; n[1] = B[4]; n[0] = ck;
    if (m.dfa.b58_mod != 0) {
        // literal-suffix patterns (Cat$): the last k symbols are the payload mod 58^k. 13 multiply-adds over the 16-bit halves
        // of the payload (each product < 2^46) and one Barrett reduction decide all but one key in 58^k before any division.
        unsigned long long s = 0;
        VG_UNROLL for (int j = 0; j < 13; j++) {
            const uint32_t hw = (j == 12) ? n[6] : ((n[j >> 1] >> (16 * (j & 1))) & 0xffffu);
            s += (unsigned long long)hw * m.dfa.b58_w[j];
        }
        const unsigned long long M = m.dfa.b58_mod, mu = ((unsigned long long)m.dfa.b58_mu_hi << 32) | m.dfa.b58_mu_lo;
        const unsigned long long q = vg_umul64hi(s, mu);           // floor(s / M) - {0, 1, 2}
        uint32_t r = (uint32_t)(s - q * M);                        // < 3 M < 2^32
        if (r >= m.dfa.b58_mod) r -= m.dfa.b58_mod;
        if (r >= m.dfa.b58_mod) r -= m.dfa.b58_mod;
        bool any = false;
        VG_UNROLL for (int k = 0; k < 8; k++) any = any || ((uint32_t)k < m.dfa.b58_nvals && r == m.dfa.b58_vals[k]);
        if (!any) return false;
    }
#if VG_B58_TABLE
    if (v.b58) return vg_match_base58check_tab<VERSION>(m, v, n);
#endif
    const uint32_t zeros = vg_base58_leading_ones<VERSION>(n);
    // seven base-58^5 groups, least significant first (58^35 > 2^200)
    const uint32_t D = 656356768u;   // 58^5
    uint32_t grp[7];
    VG_UNROLL for (int r = 0; r < 7; r++) {
        uint64_t rem = 0;
        VG_UNROLL for (int i = 6; i >= 0; i--) {
            if (i <= 6 - r || r == 0) {          // quotient shrinks by >29 bits per round: limbs above 6-r are already zero
                uint64_t cur = (rem << 32) | n[i];
                uint32_t q = (uint32_t)(cur / D);
                rem = cur - (uint64_t)q * D;
                n[i] = q;
            }
        }
        grp[r] = (uint32_t)rem;
    }
    const uint32_t nc = m.dfa.n_classes;
    uint32_t st = m.dfa.start;
    const uint32_t c_one = v.sym[0];
    for (uint32_t z = 0; z < zeros; z++) st = vg_dfa_step(v, nc, st, c_one);
    bool started = false;
    VG_UNROLL for (int r = 6; r >= 0; r--) {
        uint32_t g = grp[r];
        uint32_t d4 = g / 11316496u; g -= d4 * 11316496u;     // 58^4
        uint32_t d3 = g / 195112u;   g -= d3 * 195112u;       // 58^3
        uint32_t d2 = g / 3364u;     g -= d2 * 3364u;         // 58^2
        uint32_t d1 = g / 58u;       uint32_t d0 = g - d1 * 58u;
        uint32_t ds[5] = {d4, d3, d2, d1, d0};
        VG_UNROLL for (int k = 0; k < 5; k++) {
            started = started || ds[k] != 0;
            if (started) st = vg_dfa_step(v, nc, st, v.sym[ds[k]]);
        }
    }
    return vg_dfa_accepts(v, st);
}

// ---- Bech32 / Bech32m (BIP173 / BIP350) ----
VG_DEV uint32_t vg_bech32_polymod_step(uint32_t chk, uint32_t v) {
    uint32_t b = chk >> 25;
    chk = ((chk & 0x1ffffffu) << 5) ^ v;
    chk ^= (0u - ((b >> 0) & 1u)) & 0x3b6a57b2u;
    chk ^= (0u - ((b >> 1) & 1u)) & 0x26508e6du;
    chk ^= (0u - ((b >> 2) & 1u)) & 0x1ea119fau;
    chk ^= (0u - ((b >> 3) & 1u)) & 0x3d4233ddu;
    chk ^= (0u - ((b >> 4) & 1u)) & 0x2a1462b3u;
    return chk;
}
// The polymod step is chk -> L(chk) ^ v with L linear over GF(2), so the checksum of NSYM data symbols is
//   L^(NSYM+6)(c0) ^ XOR_j L^(NSYM+5-j)(v_j)            (c0 = state after the hrp/version prefix, 6 zero steps at the end).
// tab[32*j + v] = L^(NSYM+5-j)(v), tab[32*NSYM] = L^(NSYM+6)(c0) ^ final constant: one shared-memory load and one XOR per
// symbol instead of the 12-instruction step. A row is 32 consecutive words = 32 banks, so any mix of symbols in a warp is
// conflict-free.
VG_DEV uint32_t vg_bech32_shift(uint32_t v, uint32_t n) {
    for (uint32_t i = 0; i < n; i++) v = vg_bech32_polymod_step(v, 0);
    return v;
}
template <int NSYM>
VG_DEV void vg_bech32_fill_table(uint32_t *tab, uint32_t witver, uint32_t final_xor, uint32_t tid, uint32_t nthreads) {
    for (uint32_t e = tid; e < (uint32_t)NSYM * 32u; e += nthreads) tab[e] = vg_bech32_shift(e & 31u, (uint32_t)NSYM + 5u - (e >> 5));
    if (tid == 0) {
        const uint32_t pre[6] = {3, 3, 0, 2, 3, witver};     // hrp expand of "bc", then the witness version
        uint32_t chk = 1;
        for (int i = 0; i < 6; i++) chk = vg_bech32_polymod_step(chk, pre[i]);
        tab[NSYM * 32] = vg_bech32_shift(chk, (uint32_t)NSYM + 6u) ^ final_xor;
    }
}
#define VG_BECH32_CONST 1u
#define VG_BECH32M_CONST 0x2bc830a3u

// Segwit address of an NSYM-symbol program: the 4-character head ("bc1q" / "bc1p") is pre-consumed by the host
// (dfa.start), then NSYM data symbols and 6 checksum symbols (BIP173 / BIP350). B = program bytes as big-endian words
// (+1 zero word).
template <int NSYM>
VG_DEV bool vg_match_segwit(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t *B, uint32_t witver, uint32_t final_xor) {
    const uint32_t nc = m.dfa.n_classes;
    uint32_t st = m.dfa.start;
    uint32_t chk;
    if (v.poly && m.dfa.chk_mask) {
        // literal-suffix patterns: the checksum alone decides all but one key in 32^k; only those run the DFA
        chk = v.poly[NSYM * 32];
        VG_UNROLL for (int j = 0; j < NSYM; j++) {
            const int bit = 5 * j, wi = bit >> 5, off = bit & 31;
            uint32_t val = off <= 27 ? (B[wi] >> (27 - off)) & 31u : ((B[wi] << (off - 27)) | (B[wi + 1] >> (59 - off))) & 31u;
            chk ^= v.poly[32 * j + val];
        }
        if ((chk & m.dfa.chk_mask) != m.dfa.chk_value) return false;
#ifndef VG_HOST_EMU
        #pragma unroll 1
#endif
        for (int j = 0; j < NSYM; j++) {
            const uint32_t bit = 5u * (uint32_t)j, wi = bit >> 5, off = bit & 31u;
            uint32_t w0 = 0, w1 = 0;
            VG_UNROLL for (int i = 0; i < ((NSYM - 1) * 5) / 32 + 1; i++) { w0 = (wi == (uint32_t)i) ? B[i] : w0; w1 = (wi == (uint32_t)i) ? B[i + 1] : w1; }
            unsigned long long two = ((unsigned long long)w0 << 32) | w1;
            st = vg_dfa_step(v, nc, st, v.sym[(uint32_t)(two >> (59u - off)) & 31u]);
        }
    } else if (v.poly) {
        chk = v.poly[NSYM * 32];
        VG_UNROLL for (int j = 0; j < NSYM; j++) {
            const int bit = 5 * j, wi = bit >> 5, off = bit & 31;
            uint32_t val = off <= 27 ? (B[wi] >> (27 - off)) & 31u : ((B[wi] << (off - 27)) | (B[wi + 1] >> (59 - off))) & 31u;
            chk ^= v.poly[32 * j + val];
            st = vg_dfa_step(v, nc, st, v.sym[val]);
        }
    } else {
        chk = 1;
        const uint32_t pre[6] = {3, 3, 0, 2, 3, witver};
        VG_UNROLL for (int i = 0; i < 6; i++) chk = vg_bech32_polymod_step(chk, pre[i]);
        VG_UNROLL for (int j = 0; j < NSYM; j++) {
            const int bit = 5 * j, wi = bit >> 5, off = bit & 31;
            uint32_t val = off <= 27 ? (B[wi] >> (27 - off)) & 31u : ((B[wi] << (off - 27)) | (B[wi + 1] >> (59 - off))) & 31u;
            chk = vg_bech32_polymod_step(chk, val);
            st = vg_dfa_step(v, nc, st, v.sym[val]);
        }
        VG_UNROLL for (int i = 0; i < 6; i++) chk = vg_bech32_polymod_step(chk, 0);
        chk ^= final_xor;
    }
    VG_UNROLL for (int i = 0; i < 6; i++) st = vg_dfa_step(v, nc, st, v.sym[(chk >> (5 * (5 - i))) & 31u]);
    return vg_dfa_accepts(v, st);
}
// P2WPKH: Bech32, witness v0, 20-byte program (32 symbols). dg = the program bytes as little-endian words.
VG_DEV bool vg_match_p2wpkh(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    uint32_t B[6];
    VG_UNROLL for (int i = 0; i < 5; i++) B[i] = vg_bswap(dg[i]);
    B[5] = 0;
    return vg_match_segwit<32>(m, v, B, 0u, VG_BECH32_CONST);
}
// P2TR: Bech32m, witness v1, 32-byte program (256 bits + 4 zero pad bits = 52 symbols).
VG_DEV bool vg_match_p2tr(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    uint32_t B[9];
    VG_UNROLL for (int i = 0; i < 8; i++) B[i] = vg_bswap(dg[i]);
    B[8] = 0;
    return vg_match_segwit<52>(m, v, B, 1u, VG_BECH32M_CONST);
}

// Ethereum: "0x" pre-consumed by the host; 40 hex symbols in EIP-55 case (symbol = nibble, +16 when upper-cased).
// When the DFA does not distinguish a-f from A-F (m.dfa.flags bit 0 clear) the checksum Keccak is skipped.
VG_DEV bool vg_match_eth(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    const uint32_t nc = m.dfa.n_classes;
    uint32_t st = m.dfa.start;
    uint32_t hsh[5] = {0, 0, 0, 0, 0};
    const bool cased = (m.dfa.flags & 1u) != 0;
    // literal hex suffix: the last digest bits decide all but one key in 16^k
    if (m.dfa.chk_mask != 0 && (vg_bswap(dg[4]) & m.dfa.chk_mask) != m.dfa.chk_value) return false;
    if (cased && m.dfa.fold_table != 0) {
        // the case-folded pattern on the lower-case hex: a necessary condition that needs no checksum hash
        uint32_t sf = m.dfa.fold_start;
        VG_UNROLL for (int i = 0; i < 20; i++) {
            uint32_t b = (dg[i >> 2] >> (8 * (i & 3))) & 0xffu;
            sf = v.table[m.dfa.fold_table + sf * m.dfa.fold_classes + m.dfa.fold_sym[b >> 4]];
            sf = v.table[m.dfa.fold_table + sf * m.dfa.fold_classes + m.dfa.fold_sym[b & 15u]];
        }
        if (v.accept[m.dfa.fold_accept + sf] == 0) return false;
    }
    if (cased) {
        uint32_t msg[10];
        VG_UNROLL for (int w = 0; w < 5; w++) {
            // digest bytes 4w..4w+3 -> 8 ASCII hex chars -> two little-endian message words
            uint32_t x = dg[w];
            VG_UNROLL for (int hlf = 0; hlf < 2; hlf++) {
                uint32_t b0 = (x >> (16 * hlf)) & 0xffu, b1 = (x >> (16 * hlf + 8)) & 0xffu;
                uint32_t n0 = b0 >> 4, n1 = b0 & 15u, n2 = b1 >> 4, n3 = b1 & 15u;
                uint32_t c0 = n0 + (n0 < 10 ? 48u : 87u), c1 = n1 + (n1 < 10 ? 48u : 87u);
                uint32_t c2 = n2 + (n2 < 10 ? 48u : 87u), c3 = n3 + (n3 < 10 ? 48u : 87u);
                msg[2 * w + hlf] = c0 | (c1 << 8) | (c2 << 16) | (c3 << 24);
            }
        }
        keccak_hex40(hsh, msg);
    }
    VG_UNROLL for (int i = 0; i < 20; i++) {
        uint32_t b = (dg[i >> 2] >> (8 * (i & 3))) & 0xffu;
        uint32_t hb = (hsh[i >> 2] >> (8 * (i & 3))) & 0xffu;
        uint32_t n0 = b >> 4, n1 = b & 15u;
        uint32_t s0 = n0 + ((n0 >= 10 && (hb & 0x80u)) ? 16u : 0u);
        uint32_t s1 = n1 + ((n1 >= 10 && (hb & 0x08u)) ? 16u : 0u);
        st = vg_dfa_step(v, nc, st, v.sym[s0]);
        st = vg_dfa_step(v, nc, st, v.sym[s1]);
    }
    return vg_dfa_accepts(v, st);
}

template <int FMT>
VG_DEV bool vg_confirm(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    if (FMT == VG_FMT_P2PKH || FMT == VG_FMT_P2PKH_UNCOMPRESSED) return vg_match_base58check<0>(m, v, dg);
    if (FMT == VG_FMT_P2SH_P2WPKH) return vg_match_base58check<5>(m, v, dg);
    if (FMT == VG_FMT_P2WPKH) return vg_match_p2wpkh(m, v, dg);
    if (FMT == VG_FMT_ETHEREUM) return vg_match_eth(m, v, dg);
    return vg_match_p2tr(m, v, dg);
}
