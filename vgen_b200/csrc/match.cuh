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
#else
#define VG_CLZ(x) ((uint32_t)__clz((int)(x)))
#endif

struct vg_dfa_view {
    const uint16_t *table;     // [n_states][n_classes], shared memory when it fits, else global
    const uint8_t *accept;     // [n_states]
    const uint8_t *sym;        // [64] alphabet index -> class (vg_dfa_dev::sym_class)
};

VG_DEV uint32_t vg_dfa_step(const vg_dfa_view &v, uint32_t n_classes, uint32_t st, uint32_t cls) {
    return v.table[st * n_classes + cls];
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

// Base58Check(version || h160): runs the DFA over the address symbols. dg = h160 as little-endian words.
template <int VERSION>
VG_DEV bool vg_match_base58check(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    uint32_t B[5];
    VG_UNROLL for (int i = 0; i < 5; i++) B[i] = vg_bswap(dg[i]);
    // checksum = first 4 bytes of SHA256(SHA256(version || h160))
    uint32_t msg[15];
    msg[0] = ((uint32_t)VERSION << 24) | (B[0] >> 8);
    VG_UNROLL for (int i = 1; i < 5; i++) msg[i] = (B[i - 1] << 24) | (B[i] >> 8);
    msg[5] = B[4] << 24;
    VG_UNROLL for (int i = 6; i < 15; i++) msg[i] = 0;
    uint32_t h1[8];
    sha256_short<21>(h1, msg);
    uint32_t ck = sha256_32_first_word(h1);
    // 25-byte payload as a 200-bit big-endian number in 7 limbs, n[6] = top 8 bits = version
    uint32_t n[7];
    n[6] = (uint32_t)VERSION;
    n[5] = B[0]; n[4] = B[1]; n[3] = B[2]; n[2] = B[3]; n[1] = B[4]; n[0] = ck;
    // leading zero BYTES of the payload become leading '1' symbols (alphabet index 0)
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
    return v.accept[st] != 0;
}

// Bech32 segwit v0 address of a 20-byte program: "bc1q" is pre-consumed by the host (dfa.start), then 32 data
// symbols and 6 checksum symbols (BIP173). Polymod state after hrp-expand("bc") and witness version 0 is a constant.
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
VG_DEV bool vg_match_p2wpkh(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    uint32_t B[6];
    VG_UNROLL for (int i = 0; i < 5; i++) B[i] = vg_bswap(dg[i]);
    B[5] = 0;
    uint32_t chk = 1;
    {   // hrp expand of "bc" = 3,3,0,2,3 then witness version 0
        const uint32_t pre[6] = {3, 3, 0, 2, 3, 0};
        VG_UNROLL for (int i = 0; i < 6; i++) chk = vg_bech32_polymod_step(chk, pre[i]);
    }
    const uint32_t nc = m.dfa.n_classes;
    uint32_t st = m.dfa.start;
    VG_UNROLL for (int j = 0; j < 32; j++) {
        const int bit = 5 * j, wi = bit >> 5, off = bit & 31;
        uint32_t val = off <= 27 ? (B[wi] >> (27 - off)) & 31u : ((B[wi] << (off - 27)) | (B[wi + 1] >> (59 - off))) & 31u;
        chk = vg_bech32_polymod_step(chk, val);
        st = vg_dfa_step(v, nc, st, v.sym[val]);
    }
    VG_UNROLL for (int i = 0; i < 6; i++) chk = vg_bech32_polymod_step(chk, 0);
    chk ^= 1u;
    VG_UNROLL for (int i = 0; i < 6; i++) st = vg_dfa_step(v, nc, st, v.sym[(chk >> (5 * (5 - i))) & 31u]);
    return v.accept[st] != 0;
}

// Bech32m segwit v1 address of a 32-byte program (P2TR, BIP350): "bc1p" is pre-consumed by the host, then 52 data symbols
// (256 bits + 4 zero pad bits) and 6 checksum symbols. dg = the program bytes as little-endian words.
VG_DEV bool vg_match_p2tr(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    uint32_t B[9];
    VG_UNROLL for (int i = 0; i < 8; i++) B[i] = vg_bswap(dg[i]);
    B[8] = 0;
    uint32_t chk = 1;
    {   // hrp expand of "bc" = 3,3,0,2,3 then witness version 1
        const uint32_t pre[6] = {3, 3, 0, 2, 3, 1};
        VG_UNROLL for (int i = 0; i < 6; i++) chk = vg_bech32_polymod_step(chk, pre[i]);
    }
    const uint32_t nc = m.dfa.n_classes;
    uint32_t st = m.dfa.start;
    VG_UNROLL for (int j = 0; j < 52; j++) {
        const int bit = 5 * j, wi = bit >> 5, off = bit & 31;
        uint32_t val = off <= 27 ? (B[wi] >> (27 - off)) & 31u : ((B[wi] << (off - 27)) | (B[wi + 1] >> (59 - off))) & 31u;
        chk = vg_bech32_polymod_step(chk, val);
        st = vg_dfa_step(v, nc, st, v.sym[val]);
    }
    VG_UNROLL for (int i = 0; i < 6; i++) chk = vg_bech32_polymod_step(chk, 0);
    chk ^= 0x2bc830a3u;
    VG_UNROLL for (int i = 0; i < 6; i++) st = vg_dfa_step(v, nc, st, v.sym[(chk >> (5 * (5 - i))) & 31u]);
    return v.accept[st] != 0;
}

// Ethereum: "0x" pre-consumed by the host; 40 hex symbols in EIP-55 case (symbol = nibble, +16 when upper-cased).
// When the DFA does not distinguish a-f from A-F (m.dfa.flags bit 0 clear) the checksum Keccak is skipped.
VG_DEV bool vg_match_eth(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    const uint32_t nc = m.dfa.n_classes;
    uint32_t st = m.dfa.start;
    uint32_t hsh[5] = {0, 0, 0, 0, 0};
    const bool cased = (m.dfa.flags & 1u) != 0;
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
    return v.accept[st] != 0;
}

template <int FMT>
VG_DEV bool vg_confirm(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]) {
    if (FMT == VG_FMT_P2PKH || FMT == VG_FMT_P2PKH_UNCOMPRESSED) return vg_match_base58check<0>(m, v, dg);
    if (FMT == VG_FMT_P2SH_P2WPKH) return vg_match_base58check<5>(m, v, dg);
    if (FMT == VG_FMT_P2WPKH) return vg_match_p2wpkh(m, v, dg);
    if (FMT == VG_FMT_ETHEREUM) return vg_match_eth(m, v, dg);
    return vg_match_p2tr(m, v, dg);
}
