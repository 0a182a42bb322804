// Integer issue-rate microbenchmarks for the roofline denominator: pure IMAD,
// pure IADD3/LOP3/SHF, IMAD.WIDE (mad.wide.u32 and carry-chained), and 1:1 mixes, on all SMs.
// Prints one JSON object; bench.py stores it under profiles/.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
#define UNROLL 16

template <int KIND>
__global__ void __launch_bounds__(256) k(uint32_t *out, uint32_t seed) {
    uint32_t a0 = seed + threadIdx.x, a1 = a0 * 3 + 1, a2 = a0 * 5 + 2, a3 = a0 * 7 + 3;
    uint32_t b0 = a0 ^ 0x9e3779b9u, b1 = a1 ^ 0x7f4a7c15u, b2 = a2 ^ 0x85ebca6bu, b3 = a3 ^ 0xc2b2ae35u;
    unsigned long long w0 = a0, w1 = a1, w2 = a2, w3 = a3;
    for (int i = 0; i < ITERS; i++) {
#pragma unroll
        for (int u = 0; u < UNROLL; u++) {
            if (KIND == 0) {          // IMAD (32-bit mad.lo)
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a0) : "r"(b0), "r"(b1));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a1) : "r"(b1), "r"(b2));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a2) : "r"(b2), "r"(b3));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a3) : "r"(b3), "r"(b0));
            } else if (KIND == 1) {   // ALU: LOP3 + SHF + IADD3 mix
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a0) : "r"(b0), "r"(b1));
                asm volatile("shf.l.wrap.b32 %0,%0,%0,%1;" : "+r"(a1) : "r"(7));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a2) : "r"(b2));
                asm volatile("lop3.b32 %0,%0,%1,%2,0xe8;" : "+r"(a3) : "r"(b3), "r"(b0));
            } else if (KIND == 2) {   // IMAD.WIDE (mad.wide.u32 with 64-bit accumulate)
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w0) : "r"(b0), "r"(b1));
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w1) : "r"(b1), "r"(b2));
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w2) : "r"(b2), "r"(b3));
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w3) : "r"(b3), "r"(b0));
            } else if (KIND == 3) {   // carry-chained pair (mad.lo.cc + madc.hi.cc -> IMAD.WIDE.U32.X)
                asm volatile("mad.lo.cc.u32 %0,%2,%3,%0; madc.hi.cc.u32 %1,%2,%3,%1;" : "+r"(a0), "+r"(a1) : "r"(b0), "r"(b1));
                asm volatile("madc.lo.cc.u32 %0,%2,%3,%0; madc.hi.cc.u32 %1,%2,%3,%1;" : "+r"(a2), "+r"(a3) : "r"(b2), "r"(b3));
                asm volatile("madc.lo.cc.u32 %0,%2,%3,%0; madc.hi.cc.u32 %1,%2,%3,%1;" : "+r"(b0), "+r"(b1) : "r"(a0), "r"(a1));
                asm volatile("madc.lo.cc.u32 %0,%2,%3,%0; madc.hi.u32 %1,%2,%3,%1;" : "+r"(b2), "+r"(b3) : "r"(a2), "r"(a3));
            } else if (KIND == 4) {   // 1:1 IMAD : ALU
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a0) : "r"(b0), "r"(b1));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a1) : "r"(b1), "r"(b2));
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a2) : "r"(b2), "r"(b3));
                asm volatile("shf.l.wrap.b32 %0,%0,%0,%1;" : "+r"(a3) : "r"(11));
            } else if (KIND == 5) {   // 1:1 IMAD.WIDE : ALU
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w0) : "r"(b0), "r"(b1));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a1) : "r"(b1), "r"(b2));
                asm volatile("mad.wide.u32 %0,%1,%2,%0;" : "+l"(w2) : "r"(b2), "r"(b3));
                asm volatile("shf.l.wrap.b32 %0,%0,%0,%1;" : "+r"(a3) : "r"(11));
            } else {                  // 1:3 IMAD : ALU (the hash-heavy mix of the scan kernel)
                asm volatile("mad.lo.u32 %0,%0,%1,%2;" : "+r"(a0) : "r"(b0), "r"(b1));
                asm volatile("lop3.b32 %0,%0,%1,%2,0x96;" : "+r"(a1) : "r"(b1), "r"(b2));
                asm volatile("add.u32 %0,%0,%1;" : "+r"(a2) : "r"(b2));
                asm volatile("shf.l.wrap.b32 %0,%0,%0,%1;" : "+r"(a3) : "r"(11));
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
    double inst = (double)blocks * 256 * (double)ITERS * UNROLL * 4;   // thread-level instructions (4 per unrolled step)
    return inst / (best * 1e-3) / 1e12;   // T thread-instructions / s
}

int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    uint32_t *d; cudaMalloc(&d, (size_t)sms * 8 * 256 * 4);
    int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz_max\": %d, \"unit\": \"T thread-inst/s\"", p.name, sms, clk);
    printf(", \"imad\": %.3f", run<0>(sms, d, 8));
    printf(", \"alu_lop3_shf_iadd\": %.3f", run<1>(sms, d, 8));
    printf(", \"imad_wide\": %.3f", run<2>(sms, d, 8));
    printf(", \"imad_wide_carry_chain\": %.3f", run<3>(sms, d, 8));
    printf(", \"mix_imad_alu_1_1\": %.3f", run<4>(sms, d, 8));
    printf(", \"mix_imadwide_alu_1_1\": %.3f", run<5>(sms, d, 8));
    printf(", \"mix_imad_alu_1_3\": %.3f", run<6>(sms, d, 8));
    printf(", \"nominal_64_lanes_at_max_clock\": %.3f}\n", (double)sms * 64 * clk * 1e3 / 1e12);
    cudaFree(d);
    return 0;
}
