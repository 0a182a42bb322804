// Integer issue-rate microbenchmarks for the roofline denominator on all SMs: the multiplier (FMA-heavy) pipe
// (IMAD, IMAD.WIDE, the carry-chained IMAD.WIDE.X of fe.cuh, IMAD.IADD), the ALU pipe (LOP3/SHF, IADD3) and mixes.
// Prints one JSON object; bench.py runs it inside the bench run and embeds the numbers.
//
// Every kernel is written so that ptxas cannot shrink it, and tests/test_peak_bench_sass.py checks the SASS histogram of each
// kernel against the instruction count used below. Round 1's version was wrong in two places (profiles/peak_int.json, r01):
//   * `add a, a, b` twice in a row is fused into ONE IADD3 (a + b + b), and a lone 2-input add is emitted as IMAD.IADD on the
//     multiplier pipe, so its "pure ALU 19.3 T" was neither pure nor 19.3 (16.9 executed);
//   * `mad.wide a*b + acc` with loop-invariant a, b was hoisted: the loop ran IADD3 + IADD3.X, not IMAD.WIDE.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
#define UNROLL 16

#define LOP(x, y, z) asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(x) : "r"(y), "r"(z))
#define SHF(x, n) asm volatile("shf.l.wrap.b32 %0,%0,%0,%1;" : "+r"(x) : "r"(n))
#define MAD(x, y, z) asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(x) : "r"(y), "r"(z))
// 3-input add: one IADD3 (ALU pipe); written as two PTX adds, which ptxas fuses
#define ADD3(x, y, z) asm volatile("{ .reg .u32 t; add.u32 t,%0,%1; add.u32 %0,t,%2; }" : "+r"(x) : "r"(y), "r"(z))

enum { K_IMAD, K_ALU_LOGIC, K_ALU_IADD3, K_IMAD_WIDE, K_IMAD_WIDE_X, K_IMAD_IADD, K_MIX_1_1, K_MIX_1_3, K_MIX_WIDEX_ALU, K_COUNT };

template <int KIND>
__global__ void __launch_bounds__(256) k(uint32_t *out, uint32_t seed) {
    uint32_t a0 = seed + threadIdx.x, a1 = a0 * 3 + 1, a2 = a0 * 5 + 2, a3 = a0 * 7 + 3;
    uint32_t b0 = a0 ^ 0x9e3779b9u, b1 = a1 ^ 0x7f4a7c15u, b2 = a2 ^ 0x85ebca6bu, b3 = a3 ^ 0xc2b2ae35u;
    unsigned long long w0 = a0, w1 = a1, w2 = a2, w3 = a3;
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            if (KIND == K_IMAD) {                 // 4 x IMAD
                MAD(a0, b0, b1); MAD(a1, b1, b2); MAD(a2, b2, b3); MAD(a3, b3, b0);
            } else if (KIND == K_ALU_LOGIC) {     // 2 x LOP3 + 2 x SHF: stays on the ALU pipe
                LOP(a0, b0, b1); SHF(a1, 7); LOP(a2, b2, b3); SHF(a3, 13);
            } else if (KIND == K_ALU_IADD3) {     // 4 x IADD3 (3-input adds cannot become IMAD.IADD)
                ADD3(a0, b0, b1); ADD3(a1, b1, b2); ADD3(a2, b2, b3); ADD3(a3, b3, b0);
            } else if (KIND == K_IMAD_WIDE) {     // 4 x mad.wide.u32 with a 64-bit addend, one factor changing every step. ptxas emits a
                                                  // mix: IMAD.WIDE.U32 with the addend fused, or IMAD.WIDE.U32 + IADD3 + IADD3.X (4 SASS per step)
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w0) : "r"((uint32_t)w1), "r"(b0));
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w1) : "r"((uint32_t)w2), "r"(b1));
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w2) : "r"((uint32_t)w3), "r"(b2));
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w3) : "r"((uint32_t)w0), "r"(b3));
            } else if (KIND == K_IMAD_WIDE_X) {   // carry-chained pairs (mad.lo.cc + madc.hi.cc -> IMAD.WIDE.U32.X), as in fe_mul_wide
                asm volatile("mad.lo.cc.u32 %0,%2,%3,%0; madc.hi.cc.u32 %1,%2,%3,%1;" : "+r"(a0), "+r"(a1) : "r"(b0), "r"(b1));
                asm volatile("madc.lo.cc.u32 %0,%2,%3,%0; madc.hi.cc.u32 %1,%2,%3,%1;" : "+r"(a2), "+r"(a3) : "r"(b2), "r"(b3));
                asm volatile("madc.lo.cc.u32 %0,%2,%3,%0; madc.hi.cc.u32 %1,%2,%3,%1;" : "+r"(b0), "+r"(b1) : "r"(a0), "r"(a1));
                asm volatile("madc.lo.cc.u32 %0,%2,%3,%0; madc.hi.u32 %1,%2,%3,%1;" : "+r"(b2), "+r"(b3) : "r"(a2), "r"(a3));
            } else if (KIND == K_IMAD_IADD) {     // 2-input adds: ptxas itself spreads them over both pipes, IMAD.IADD : IADD3 = 1 : 1
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a0) : "r"(a1));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a1) : "r"(a2));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a2) : "r"(a3));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a3) : "r"(a0));
            } else if (KIND == K_MIX_1_1) {       // 2 x IMAD + LOP3 + SHF
                MAD(a0, b0, b1); LOP(a1, b1, b2); MAD(a2, b2, b3); SHF(a3, 11);
            } else if (KIND == K_MIX_1_3) {       // IMAD + LOP3 + IADD3 + SHF: the hash-heavy mix of the scan kernel
                MAD(a0, b0, b1); LOP(a1, b1, b2); ADD3(a2, b2, b3); SHF(a3, 11);
            } else {                              // IMAD.WIDE.X pair chain + 2 ALU: the field-multiplication mix
                asm volatile("mad.lo.cc.u32 %0,%2,%3,%0; madc.hi.cc.u32 %1,%2,%3,%1;" : "+r"(a0), "+r"(a1) : "r"(b0), "r"(b1));
                asm volatile("madc.lo.cc.u32 %0,%2,%3,%0; madc.hi.u32 %1,%2,%3,%1;" : "+r"(b0), "+r"(b1) : "r"(a0), "r"(a1));
                LOP(a2, b2, b3); SHF(a3, 11);
            }
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0 ^ a1 ^ a2 ^ a3 ^ b0 ^ b1 ^ b2 ^ b3 ^ (uint32_t)(w0 ^ w1 ^ w2 ^ w3) ^ (uint32_t)((w0 ^ w1 ^ w2 ^ w3) >> 32);
}

template <int KIND>
double run(int sms, uint32_t *d, int blocks_per_sm) {
    int blocks = sms * blocks_per_sm;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<KIND><<<blocks, 256>>>(d, 1); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0); k<KIND><<<blocks, 256>>>(d, r); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
    }
    double inst = (double)blocks * 256 * (double)ITERS * UNROLL * 4;   // thread-level SASS instructions (4 per unrolled step, checked by the SASS test)
    return inst / (best * 1e-3) / 1e12;   // T thread-instructions / s
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    uint32_t *d; cudaMalloc(&d, (size_t)sms * 8 * 256 * 4);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz_max\": %d, \"unit\": \"T thread-inst/s (SASS instructions)\"", p.name, sms, clk);
    printf(", \"imad\": %.3f", run<K_IMAD>(sms, d, 8));
    printf(", \"alu_lop3_shf\": %.3f", run<K_ALU_LOGIC>(sms, d, 8));
    printf(", \"alu_iadd3\": %.3f", run<K_ALU_IADD3>(sms, d, 8));
    printf(", \"mad_wide_u32_as_compiled\": %.3f", run<K_IMAD_WIDE>(sms, d, 8));
    printf(", \"imad_wide_carry_chain\": %.3f", run<K_IMAD_WIDE_X>(sms, d, 8));
    printf(", \"add_u32_as_compiled\": %.3f", run<K_IMAD_IADD>(sms, d, 8));
    printf(", \"mix_imad_alu_1_1\": %.3f", run<K_MIX_1_1>(sms, d, 8));
    printf(", \"mix_imad_alu_1_3\": %.3f", run<K_MIX_1_3>(sms, d, 8));
    printf(", \"mix_imadwidex_alu_1_1\": %.3f", run<K_MIX_WIDEX_ALU>(sms, d, 8));
    printf(", \"nominal_64_lanes_at_max_clock\": %.3f}\n", (double)sms * 64 * clk * 1e3 / 1e12);
    cudaFree(d);
    return 0;
}
