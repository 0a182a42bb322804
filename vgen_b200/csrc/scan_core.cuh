// Per-thread body of the scan ("walk") kernel, written so that the identical code runs under the
// host emulation (VG_HOST_EMU) in tests/ and inside the __global__ wrappers in scan_kernels.cu.
//
// One launch covers a contiguous window of W = n_threads * 2*half keys. Thread t owns the 2*half keys
// [t*2H, (t+1)*2H) of the window, centred on C_t = (window_base + t*2H + H) * G, and derives them as
// C_t +/- i*G (i = 1..H) with ONE field inversion for the whole batch (Montgomery's trick; the +/-
// pair shares its inverse because x(i*G) = x(-i*G)). The same inversion also yields C_t + W*G, the
// centre for the next window. This replaces the reference's `main` entry point
// (src/shaders/search.wgsl:2-31: one Jacobian add + one 505-multiplication Fermat inversion per key)
// and its host loop's per-key address/regex work (src/gpu.rs:1030-1093): digests are matched on the
// device and only hits leave the chip.
#pragma once
#include "ec.cuh"
#include "hash.cuh"
#include "scan_types.h"

// Device-side bounds checks of the checked build (-DVG_CHECKED): a failed check records its code in P.violations and the host
// turns it into an error (engine.cpp harvest). Compiled out of the product build. Codes: 1 thread id, 2 scratch index, 3 table
// index, 4 hit slot, 5 degenerate-list slot, 6 point-buffer index, 7 DFA table index, 8 DFA accept index, 9 shared-memory size,
// 10 fixed-base table index, 11 dump index.
#if defined(VG_CHECKED) && !defined(VG_HOST_EMU)
#define VG_ASSERT_AT(viol, cond, code) do { if (!(cond) && (viol)) { atomicCAS((viol), 0u, (uint32_t)(code)); atomicAdd((viol) + 1, 1u); } } while (0)
#else
#define VG_ASSERT_AT(viol, cond, code) ((void)0)
#endif
#define VG_ASSERT(P, cond, code) VG_ASSERT_AT((P).violations, cond, code)

// VG_TAP_SCALAR = 1 (P2TR): the output key is computed from the key's SCALAR, Q = (t + k) * G for even y(k*G) and (t + n - k) * G for
// odd, entirely from the fixed-base table: one table addition fewer per key than P + t*G (the point P itself is only hashed).
#ifndef VG_TAP_SCALAR
#define VG_TAP_SCALAR 1
#endif

// digest words of N points at once (T = uint32_t or vw<N>); xv / yv = little-endian limbs of canonical coordinates
template <int FMT, class T>
VG_DEV void vg_digest_lanes(T dg[VG_DIGEST_WORDS], const T xv[8], const T yv[8]) {
    VG_UNROLL for (int i = 5; i < VG_DIGEST_WORDS; i++) vg_splat(dg[i], 0u);
    if (FMT == VG_FMT_P2PKH || FMT == VG_FMT_P2WPKH) {
        hash160_pub33(dg, (yv[0] & 1u) + 2u, xv);
    } else if (FMT == VG_FMT_P2PKH_UNCOMPRESSED) {
        hash160_pub65(dg, xv, yv);
    } else if (FMT == VG_FMT_P2SH_P2WPKH) {
        T h[5]; hash160_pub33(h, (yv[0] & 1u) + 2u, xv); hash160_p2sh_script(dg, h);
    } else {   // P2TR fast path yields the even-Y internal key x (as src/shaders/search_p2tr.wgsl:104-113)
        VG_UNROLL for (int i = 0; i < 8; i++) dg[i] = vg_bswap(xv[7 - i]);
    }
}
template <int FMT>
VG_DEV void vg_digest_of_point(uint32_t dg[VG_DIGEST_WORDS], const fe &x, const fe &y) {
    if (FMT == VG_FMT_ETHEREUM) {
        VG_UNROLL for (int i = 5; i < VG_DIGEST_WORDS; i++) dg[i] = 0;
        keccak_eth_address(dg, x, y);
    } else {
        vg_digest_lanes<FMT, uint32_t>(dg, x.v, y.v);
    }
}
// lanes hashed together by one thread. Measured on B200 (r01): ILP 2 costs 128 registers and runs 7 % slower than ILP 1
// at 118 registers (5.41 vs 5.82 Gkeys/s), so 1 is the default; Keccak is always 1 (register budget).
#ifndef VG_ILP
#define VG_ILP 1
#endif
template <int FMT> struct vg_ilp { static const int N = (FMT == VG_FMT_ETHEREUM) ? 1 : VG_ILP; };

VG_DEV bool vg_prefilter(const vg_matcher_dev &m, const uint32_t dg[VG_DIGEST_WORDS]) {
    if (m.kind == VG_MATCH_ALL || m.kind == VG_MATCH_DFA) return true;
    if (m.kind == VG_MATCH_EXACT) {
        uint32_t d = 0;
        VG_UNROLL for (int i = 0; i < VG_DIGEST_WORDS; i++) d |= (i < m.exact_words) ? (dg[i] ^ m.exact[i]) : 0u;   // static indices: dg stays in registers
        return d == 0;
    }
    uint32_t hi = vg_bswap(dg[0]), lo = vg_bswap(dg[1]);
    bool hit = false;
    for (int k = 0; k < m.n_intervals; k++) {
        const uint32_t *iv = m.lo_hi[k];
        bool ge = hi > iv[0] || (hi == iv[0] && lo >= iv[1]);
        bool le = hi < iv[2] || (hi == iv[2] && lo <= iv[3]);
        hit = hit || (ge && le);
    }
    return hit;
}

#ifdef VG_HOST_EMU
static inline void vg_emit_hit(const vg_scan_params &P, bool is_hit, unsigned long long idx, const uint32_t dg[VG_DIGEST_WORDS]) {
    if (!is_hit) return;
    uint32_t slot = (*P.hit_count)++;
    if (slot < P.hit_cap) {
        P.hits[slot].idx_lo = (uint32_t)idx; P.hits[slot].idx_hi = (uint32_t)(idx >> 32);
        for (int i = 0; i < VG_DIGEST_WORDS; i++) P.hits[slot].digest[i] = dg[i];
    }
}
#else
// warp-ballot compaction: one atomicAdd per warp that has hits; the count keeps growing past the
// capacity so the host can detect overflow and rescan the window in smaller pieces.
__device__ __forceinline__ void vg_emit_hit(const vg_scan_params &P, bool is_hit, unsigned long long idx, const uint32_t dg[VG_DIGEST_WORDS]) {
    unsigned mask = __ballot_sync(0xffffffffu, is_hit);
    if (mask == 0) return;
    unsigned lane = threadIdx.x & 31u, leader = __ffs(mask) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atomicAdd(P.hit_count, (uint32_t)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (is_hit) {
        uint32_t slot = base + __popc(mask & ((1u << lane) - 1u));
        if (slot < P.hit_cap) {
            VG_ASSERT(P, idx < (unsigned long long)P.n_threads * 2ull * P.half, 4);
            vg_hit_rec *h = P.hits + slot;
            h->idx_lo = (uint32_t)idx; h->idx_hi = (uint32_t)(idx >> 32);
            #pragma unroll
            for (int i = 0; i < VG_DIGEST_WORDS; i++) h->digest[i] = dg[i];
        }
    }
}
#endif

struct vg_dfa_view;
template <int FMT> VG_DEV bool vg_confirm(const vg_matcher_dev &m, const vg_dfa_view &v, const uint32_t dg[VG_DIGEST_WORDS]);

// test / dump one already-computed digest
template <int FMT, int MODE>
VG_DEV void vg_finish_point(const vg_scan_params &P, const vg_dfa_view &dv, const uint32_t dg[VG_DIGEST_WORDS], unsigned long long idx, bool valid) {
    valid = valid && idx < P.limit;
    if (MODE == VG_MODE_DUMP) {
        if (valid) {
            VG_ASSERT(P, P.dump != nullptr && idx < P.limit, 11);
            uint8_t *o = P.dump + idx * P.dump_stride;
            for (uint32_t b = 0; b < P.dump_stride; b++) {
                uint32_t wsel = 0;
                VG_UNROLL for (int i = 0; i < VG_DIGEST_WORDS; i++) wsel = ((b >> 2) == (uint32_t)i) ? dg[i] : wsel;
                o[b] = (uint8_t)(wsel >> (8 * (b & 3)));
            }
        }
    } else {
        bool hit = valid && vg_prefilter(P.m, dg);
        if (hit && (P.m.kind == VG_MATCH_DFA || P.m.confirm_dfa)) hit = vg_confirm<FMT>(P.m, dv, dg);
        vg_emit_hit(P, hit, idx, dg);
    }
}
template <int FMT, int MODE>
VG_DEV void vg_process_point(const vg_scan_params &P, const vg_dfa_view &dv, fe &x, fe &y, unsigned long long idx, bool valid) {
    fe_canon(x); fe_canon(y);
    if (FMT == VG_FMT_P2TR) {
        // stage 1 of the P2TR path: hand the public key to stage 2 (taproot.cuh); MODE is applied there.
        // VG_TAP_SCALAR: (x, y) as it is — stage 2 needs x for the tweak hash and the parity of y for the sign of the scalar;
        // else the even-Y internal key (y negated when odd), to which stage 2 adds t*G.
        valid = valid && idx < P.limit;
        VG_ASSERT(P, P.points != nullptr && idx < (unsigned long long)P.n_threads * 2ull * P.half, 6);
        uint32_t *o = P.points + idx * 16;
#if VG_TAP_SCALAR
        VG_UNROLL for (int l = 0; l < 8; l++) { o[l] = valid ? x.v[l] : 0u; o[8 + l] = valid ? y.v[l] : 0u; }
#else
        fe ny; fe_neg(ny, y); fe_canon(ny);
        const bool odd = (y.v[0] & 1u) != 0;
        VG_UNROLL for (int l = 0; l < 8; l++) { o[l] = valid ? x.v[l] : 0u; o[8 + l] = valid ? (odd ? ny.v[l] : y.v[l]) : 0u; }
#endif
        return;
    }
    uint32_t dg[VG_DIGEST_WORDS];
    vg_digest_of_point<FMT>(dg, x, y);
    vg_finish_point<FMT, MODE>(P, dv, dg, idx, valid);
}
// the +/- pair of one walk step: both digests are computed in lock-step (vw<2>) where the format allows
template <int FMT, int MODE>
VG_DEV void vg_process_pair(const vg_scan_params &P, const vg_dfa_view &dv, fe &x0, fe &y0, unsigned long long idx0,
                            fe &x1, fe &y1, unsigned long long idx1, bool valid) {
    if (vg_ilp<FMT>::N == 2) {
        fe_canon(x0); fe_canon(y0); fe_canon(x1); fe_canon(y1);
        vw<2> xv[8], yv[8], dg[VG_DIGEST_WORDS];
        VG_UNROLL for (int l = 0; l < 8; l++) { xv[l].v[0] = x0.v[l]; xv[l].v[1] = x1.v[l]; yv[l].v[0] = y0.v[l]; yv[l].v[1] = y1.v[l]; }
        vg_digest_lanes<FMT, vw<2> >(dg, xv, yv);
        uint32_t d0[VG_DIGEST_WORDS], d1[VG_DIGEST_WORDS];
        VG_UNROLL for (int i = 0; i < VG_DIGEST_WORDS; i++) { d0[i] = dg[i].v[0]; d1[i] = dg[i].v[1]; }
        vg_finish_point<FMT, MODE>(P, dv, d0, idx0, valid);
        vg_finish_point<FMT, MODE>(P, dv, d1, idx1, valid);
    } else {
        vg_process_point<FMT, MODE>(P, dv, x0, y0, idx0, valid);
        vg_process_point<FMT, MODE>(P, dv, x1, y1, idx1, valid);
    }
}

// Every VG_SYNC_ITER steps of the backward pass the CTA's warps re-align at a barrier. They share no data; the point is
// that the hot loop body is ~53 KB of straight-line code (L1.5 instruction cache: 32 KB), so warps that run the same
// stretch of it share instruction fetches, and that the warps of a CTA finish together and free its slot for the other
// lane's next window at once. Measured on B200 with 2 lanes (profiles/overlap_sync_r01.log, overlap_lockstep_r01.log):
// none 6,094, every 128 steps 6,061, 32: 6,103, 8: 6,221, 4: 6,267, 1: 6,197 Mkeys/s.
#ifndef VG_SYNC_ITER
#define VG_SYNC_ITER 4
#endif
// VG_PHASE_SPLIT = 1: the two halves of the CTA (warps 0-3 / 4-7, i.e. one warp of each half on every SM sub-partition) meet
// the barrier at DIFFERENT points of the step: the first half at the top (before the two inverse-peeling multiplications), the
// second half in the middle (after the second affine addition, before its hash). The halves then run the same code about half
// a step apart: while one half is in the multiplier-bound EC stretch (carry-chained IMAD.WIDE, half rate, ALU pipe idle) the
// other is in the ALU-bound hash stretch, and they still share instruction fetches one stretch later.
#ifndef VG_PHASE_SPLIT
#define VG_PHASE_SPLIT 0
#endif
#if !defined(VG_HOST_EMU)
#define VG_CTA_BARRIER() asm volatile("bar.sync 0;" ::: "memory")
#endif

#define VG_LOAD_FE(dst, base, stride_words) { VG_UNROLL for (int _l = 0; _l < 8; _l++) (dst).v[_l] = (base)[(size_t)_l * (stride_words)]; }
#define VG_STORE_FE(base, stride_words, src) { VG_UNROLL for (int _l = 0; _l < 8; _l++) (base)[(size_t)_l * (stride_words)] = (src).v[_l]; }

// VG_PIPELINE = 1 (hash160-of-compressed-key formats only): software pipelining inside a thread. While a point is being hashed
// (2,300 ALU-pipe instructions in a row) the field multiplications that produce the NEXT point (multiplier-pipe work) are issued
// between the hash rounds through the hook points of hash.cuh, so that a warp feeds both integer pipes at once instead of
// alternating between a multiplier-bound EC stretch and an ALU-bound hash stretch. Same results, different instruction order.
#ifndef VG_PIPELINE
#define VG_PIPELINE 0
#endif
#if VG_PIPELINE && !defined(VG_HOST_EMU)
// The next point of the walk, computed in stages at the hook points of the hash of the current one. One struct for both kinds of
// step, so that the loop body holds ONE copy of the hash (two copies, one per kind, doubled the code: 5,660 against 6,236 Mkeys/s,
// "no instruction" stall 1.28 per issue, profiles/ncu_walk_pipe_r02_summary.txt):
//   peel = false: (cx, cy) - i*G with the inverse inv_i already peeled for step i;
//   peel = true : peel the inverse of the next step (inv_i = inv * pre, inv *= dx), then (cx, cy) + i*G.
struct vg_stage_next {
    const fe &cx, &cy; const uint32_t *g; const uint32_t *pre_ptr; size_t stride; fe &inv, &inv_i; bool peel, active;
    fe lam, x3, y3;
    VG_DEVM void at(int k) {
        if (!active) return;
        if (k == 0) {
            if (peel) { fe pre; VG_LOAD_FE(pre, pre_ptr, stride); fe_mul(inv_i, inv, pre); }
        } else if (k == 2) {
            if (peel) {
                fe gx, dx; VG_UNROLL for (int l = 0; l < 8; l++) gx.v[l] = g[l];
                fe_sub(dx, gx, cx);
                if (fe_is_zero_raw(dx)) fe_set_one(dx);
                fe_mul(inv, inv, dx);
            }
        } else if (k == 4) {
            fe gy, dy; VG_UNROLL for (int l = 0; l < 8; l++) gy.v[l] = g[8 + l];
            if (!peel) fe_neg(gy, gy);
            fe_sub(dy, gy, cy);
            fe_mul(lam, dy, inv_i);
        } else if (k == 6) {
            fe gx; VG_UNROLL for (int l = 0; l < 8; l++) gx.v[l] = g[l];
            fe_sqr(x3, lam); fe_sub(x3, x3, cx); fe_sub(x3, x3, gx);
        } else if (k == 10) {
            fe t; fe_sub(t, cx, x3); fe_mul(y3, lam, t); fe_sub(y3, y3, cy);
        } else if (k == 14) { fe_canon(x3); fe_canon(y3); }
    }
};
#endif

// gtab: pointer to the i*G table (shared memory in the kernel), entry i-1 = 16 words
template <int FMT, int MODE>
VG_DEV void vg_walk_thread(const vg_scan_params &P, const vg_dfa_view &dv, const uint32_t *gtab, uint32_t tid) {
    const uint32_t T = P.n_threads, H = P.half;
    VG_ASSERT(P, tid < T, 1);
    fe cx, cy, sx, sy;
    VG_LOAD_FE(cx, P.centers + tid, T);
    VG_LOAD_FE(cy, P.centers + (size_t)8 * T + tid, T);
    VG_UNROLL for (int l = 0; l < 8; l++) { sx.v[l] = P.stride_pt[l]; sy.v[l] = P.stride_pt[8 + l]; }
    bool degenerate = fe_is_zero_raw(cx) && fe_is_zero_raw(cy);

    // forward pass: prefix products of dx_0 = Sx - Cx, dx_i = x(i*G) - Cx
    uint32_t *scr = P.scratch + tid;
    fe prod, dx;
    fe_sub(prod, sx, cx);
    if (fe_is_zero_raw(prod)) { degenerate = true; fe_set_one(prod); }
    VG_STORE_FE(scr, T, prod);
    for (uint32_t i = 1; i <= H; i++) {
        const uint32_t *g = gtab + (size_t)(i - 1) * 16;
        fe gx; VG_UNROLL for (int l = 0; l < 8; l++) gx.v[l] = g[l];
        fe_sub(dx, gx, cx);
        if (fe_is_zero_raw(dx)) { degenerate = true; fe_set_one(dx); }
        fe_mul(prod, prod, dx);
        VG_STORE_FE(scr + (size_t)i * 8 * T, T, prod);
    }
    fe inv;
    fe_inv(inv, prod);

    const unsigned long long idx_c = (unsigned long long)tid * (2ull * H) + H;
#if VG_SYNC_ITER && VG_PHASE_SPLIT && !defined(VG_HOST_EMU)
    const bool late_half = threadIdx.x >= blockDim.x / 2;
#endif
#if VG_PIPELINE && !defined(VG_HOST_EMU)
    if (FMT == VG_FMT_P2PKH || FMT == VG_FMT_P2WPKH) {
        // prologue: peel step H; the first point of the walk is the centre itself (it stands in for +H, which belongs to the next
        // thread's batch)
        fe inv_i, xa = cx, ya = cy;
        unsigned long long idx_a = idx_c;
        {
            const uint32_t *g = gtab + (size_t)(H - 1) * 16;
            fe gx, pre; VG_UNROLL for (int l = 0; l < 8; l++) gx.v[l] = g[l];
            VG_LOAD_FE(pre, scr + (size_t)(H - 1) * 8 * T, T);
            fe_mul(inv_i, inv, pre);
            fe_sub(dx, gx, cx);
            if (fe_is_zero_raw(dx)) fe_set_one(dx);
            fe_mul(inv, inv, dx);
        }
        // slot 2(H - i) hashes (i, +) [the centre for i = H] while (i, -) is computed; slot 2(H - i) + 1 hashes (i, -) while the inverse
        // of step i - 1 is peeled and (i - 1, +) is computed
        #pragma unroll 1
        for (uint32_t slot = 0; slot < 2 * H; slot++) {
            const uint32_t i = H - (slot >> 1);
            const bool second = (slot & 1u) != 0;
  #if VG_SYNC_ITER
            if (!second && (i % VG_SYNC_ITER) == 0) __syncthreads();
  #endif
            const uint32_t j = second ? i - 1 : i;              // step of the NEXT point
            const bool active = j >= 1;
            const uint32_t jj = active ? j : 1;
            vg_stage_next nx{cx, cy, gtab + (size_t)(jj - 1) * 16, scr + (size_t)(jj - 1) * 8 * T, (size_t)T, inv, inv_i, second, active};
            uint32_t dg[VG_DIGEST_WORDS];
            VG_UNROLL for (int w = 5; w < VG_DIGEST_WORDS; w++) dg[w] = 0;
            hash160_pub33<uint32_t, vg_stage_next>(dg, (ya.v[0] & 1u) + 2u, xa.v, nx);
            vg_finish_point<FMT, MODE>(P, dv, dg, idx_a, !degenerate);
            if (active) { xa = nx.x3; ya = nx.y3; idx_a = second ? idx_c + j : idx_c - j; }
        }
    } else
#endif
    // backward pass: peel one inverse per step, derive both neighbours, hash, match
    for (uint32_t i = H; i >= 1; i--) {
#if VG_SYNC_ITER && !defined(VG_HOST_EMU)
  #if VG_PHASE_SPLIT
        if ((i % VG_SYNC_ITER) == 0 && !late_half) VG_CTA_BARRIER();
  #else
        if ((i % VG_SYNC_ITER) == 0) __syncthreads();
  #endif
#endif
        const uint32_t *g = gtab + (size_t)(i - 1) * 16;
        fe gx, gy, pre, inv_i;
        VG_UNROLL for (int l = 0; l < 8; l++) { gx.v[l] = g[l]; gy.v[l] = g[8 + l]; }
        VG_LOAD_FE(pre, scr + (size_t)(i - 1) * 8 * T, T);
        fe_mul(inv_i, inv, pre);
        fe_sub(dx, gx, cx);
        if (fe_is_zero_raw(dx)) fe_set_one(dx);
        fe_mul(inv, inv, dx);
        // +i and -i share inv_i (x(iG) = x(-iG)). At i == H the + side belongs to the next thread's batch: the centre
        // itself takes its place so that every step handles exactly two keys.
        if (vg_ilp<FMT>::N == 2) {
            fe xp, yp, xm, ym, ngy;
            fe_neg(ngy, gy);
            affine_add_inv(xm, ym, cx, cy, gx, ngy, inv_i);
            unsigned long long idx_p = idx_c + i;
            if (i == H) { xp = cx; yp = cy; idx_p = idx_c; }
            else affine_add_inv(xp, yp, cx, cy, gx, gy, inv_i);
            vg_process_pair<FMT, MODE>(P, dv, xp, yp, idx_p, xm, ym, idx_c - i, !degenerate);
        } else {
#ifndef VG_HOST_EMU
            #pragma unroll 1
#endif
            for (int side = 0; side < 2; side++) {   // one side at a time keeps the live register set small
                fe x3, y3, by;
                unsigned long long idx = side ? idx_c - i : idx_c + i;
                if (side == 0 && i == H) { x3 = cx; y3 = cy; idx = idx_c; }
                else {
                    if (side) fe_neg(by, gy); else by = gy;
                    affine_add_inv(x3, y3, cx, cy, gx, by, inv_i);
                }
#if VG_SYNC_ITER && VG_PHASE_SPLIT && !defined(VG_HOST_EMU)
                if (side == 1 && (i % VG_SYNC_ITER) == 0 && late_half) VG_CTA_BARRIER();
#endif
                vg_process_point<FMT, MODE>(P, dv, x3, y3, idx, !degenerate);
            }
        }
    }
    // next centre: C + W*G, with inv = 1/(Sx - Cx)
    fe nx, ny;
    affine_add_inv(nx, ny, cx, cy, sx, sy, inv);
    fe_canon(nx); fe_canon(ny);
    if (degenerate) {
        fe_set_zero(nx); fe_set_zero(ny);
#ifdef VG_HOST_EMU
        uint32_t slot = (*P.degen_count)++;
#else
        uint32_t slot = atomicAdd(P.degen_count, 1u);
#endif
        VG_ASSERT(P, slot < P.degen_cap, 5);
        if (slot < P.degen_cap) P.degen_list[slot] = tid;
    }
    VG_STORE_FE(P.centers + tid, T, nx);
    VG_STORE_FE(P.centers + (size_t)8 * T + tid, T, ny);
}

// ---- cold-path thread bodies ----
// pow2[i] = 2^i * G for i = 0..255, built by one thread at context creation: 255 Jacobian doublings, then ONE field
// inversion for all 256 points (Montgomery's trick over the Z coordinates) instead of one per point (round 1: 25.5 ms of
// serial kernel time per lane; now ~4 k field multiplications, once per device). scratch: 256 * 32 words.
VG_DEV void vg_build_pow2(uint32_t *pow2, const uint32_t gxy[16], uint32_t *scratch) {
    gej_t p;
    VG_UNROLL for (int l = 0; l < 8; l++) { p.x.v[l] = gxy[l]; p.y.v[l] = gxy[8 + l]; }
    fe_set_one(p.z); p.inf = false;
    fe run; fe_set_one(run);
    for (int i = 0; i < 256; i++) {
        uint32_t *o = scratch + i * 32;
        fe_mul(run, run, p.z);
        for (int l = 0; l < 8; l++) { o[l] = p.x.v[l]; o[8 + l] = p.y.v[l]; o[16 + l] = p.z.v[l]; o[24 + l] = run.v[l]; }
        gej_t d; gej_double(d, p); p = d;
    }
    fe inv;
    fe_inv(inv, run);
    for (int i = 255; i >= 0; i--) {
        const uint32_t *o = scratch + i * 32;
        fe X, Y, Z, zi, zi2, zi3;
        for (int l = 0; l < 8; l++) { X.v[l] = o[l]; Y.v[l] = o[8 + l]; Z.v[l] = o[16 + l]; }
        if (i > 0) { fe pre; for (int l = 0; l < 8; l++) pre.v[l] = o[l - 8]; fe_mul(zi, inv, pre); fe_mul(inv, inv, Z); }
        else zi = inv;
        fe_sqr(zi2, zi); fe_mul(zi3, zi2, zi);
        fe_mul(X, X, zi2); fe_mul(Y, Y, zi3);
        fe_canon(X); fe_canon(Y);
        for (int l = 0; l < 8; l++) { pow2[i * 16 + l] = X.v[l]; pow2[i * 16 + 8 + l] = Y.v[l]; }
    }
}
// affine point of scalar base + mul*tid + add, stored limb-major at column tid of out[16][n]; zeros if invalid
VG_DEV void vg_point_of_scalar(uint32_t *out, uint32_t n, uint32_t col, const uint32_t k[8], const uint32_t *pow2, bool w8 = VG_W8) {
    fe x, y;
    bool ok = scalar_is_valid(k) && scalar_mul_G(x, y, k, pow2, w8);
    if (!ok) { fe_set_zero(x); fe_set_zero(y); }
    VG_STORE_FE(out + col, n, x);
    VG_STORE_FE(out + (size_t)8 * n + col, n, y);
}
// slow generic path for one key: full scalar multiplication, every edge case handled
template <int FMT, int MODE>
VG_DEV void vg_slow_key(const vg_scan_params &P, const vg_dfa_view &dv, const uint32_t k[8], unsigned long long idx, const uint32_t *pow2) {
    fe x, y;
    bool ok = scalar_is_valid(k) && scalar_mul_G(x, y, k, pow2, VG_W8);
    if (!ok) { fe_set_zero(x); fe_set_zero(y); }
    vg_process_point<FMT, MODE>(P, dv, x, y, idx, ok);
}
