// Integer-multiply pipe microbenchmarks for sm_100a: what does an IMAD.WIDE really cost?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o imad_micro imad_micro.cu ; run on the GPU box.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048

// 1: wide, 4 distinct register reads per instruction (multiplicand changes every iteration)
__global__ void k_wide_indep(uint32_t seed, unsigned long long* sink) {
    unsigned long long acc[8]; uint32_t b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = seed * (i + 1) + threadIdx.x; b[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) {
            uint32_t a = (uint32_t)acc[(i + 3) & 7];
            asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a), "r"(b[i]));
        }
    }
    unsigned long long s = 0;
    for (int i = 0; i < 8; i++) s ^= acc[i];
    if (s == 0x1234567ull) *sink = s;
}
// 2: wide, one multiplicand shared by the 8 consecutive instructions (operand reuse possible)
__global__ void k_wide_shared(uint32_t seed, unsigned long long* sink) {
    unsigned long long acc[8]; uint32_t a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = seed * (i + 1) + threadIdx.x; a[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u; }
    for (int it = 0; it < ITERS; it++) {
        uint32_t b = (uint32_t)(acc[0] >> 32) | 1u;
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(acc[i]) : "r"(a[i]), "r"(b));
    }
    unsigned long long s = 0;
    for (int i = 0; i < 8; i++) s ^= acc[i];
    if (s == 0x1234567ull) *sink = s;
}
// 3: wide with an immediate multiplicand
__global__ void k_wide_imm(uint32_t seed, unsigned long long* sink) {
    unsigned long long acc[8];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = seed * (i + 1) + threadIdx.x;
    for (int it = 0; it < ITERS; it++) {
        uint32_t b = (uint32_t)(acc[0] >> 32) | 1u;
        asm volatile("mad.wide.u32 %0, %1, 0x43e1f593, %0;" : "+l"(acc[0]) : "r"(b));
        asm volatile("mad.wide.u32 %0, %1, 0x79b97091, %0;" : "+l"(acc[1]) : "r"(b));
        asm volatile("mad.wide.u32 %0, %1, 0x2833e848, %0;" : "+l"(acc[2]) : "r"(b));
        asm volatile("mad.wide.u32 %0, %1, 0x8181585d, %0;" : "+l"(acc[3]) : "r"(b));
        asm volatile("mad.wide.u32 %0, %1, 0xb85045b6, %0;" : "+l"(acc[4]) : "r"(b));
        asm volatile("mad.wide.u32 %0, %1, 0xe131a029, %0;" : "+l"(acc[5]) : "r"(b));
        asm volatile("mad.wide.u32 %0, %1, 0x30644e72, %0;" : "+l"(acc[6]) : "r"(b));
        asm volatile("mad.wide.u32 %0, %1, 0xf0000001, %0;" : "+l"(acc[7]) : "r"(b));
    }
    unsigned long long s = 0;
    for (int i = 0; i < 8; i++) s ^= acc[i];
    if (s == 0x1234567ull) *sink = s;
}
// 4: 32-bit mad.lo, 3 distinct register reads
__global__ void k_lo32(uint32_t seed, unsigned long long* sink) {
    uint32_t acc[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = seed * (i + 1) + threadIdx.x; b[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(acc[i]) : "r"(acc[(i + 3) & 7]), "r"(b[i]));
    }
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s ^= acc[i];
    if (s == 0x1234567u) *sink = s;
}
// 5: 32-bit mad.hi
__global__ void k_hi32(uint32_t seed, unsigned long long* sink) {
    uint32_t acc[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = seed * (i + 1) + threadIdx.x; b[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("mad.hi.u32 %0, %1, %2, %0;" : "+r"(acc[i]) : "r"(acc[(i + 3) & 7] | 0x80000000u), "r"(b[i]));
    }
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s ^= acc[i];
    if (s == 0x1234567u) *sink = s;
}
// 6: carry chains as in the field multiplier: two chains of 4 lo/hi pairs with a shared multiplier bi (registers)
__global__ void k_chain_reg(uint32_t seed, unsigned long long* sink) {
    uint32_t e[8], o[8], a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { e[i] = seed * (i + 1) + threadIdx.x; o[i] = e[i] * 3u; a[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u; }
    for (int it = 0; it < ITERS; it++) {
        uint32_t bi = e[1] | 1u;
        asm volatile(
            "mad.lo.cc.u32 %0, %8, %12, %0;\n madc.hi.cc.u32 %1, %8, %12, %1;\n"
            "madc.lo.cc.u32 %2, %9, %12, %2;\n madc.hi.cc.u32 %3, %9, %12, %3;\n"
            "madc.lo.cc.u32 %4, %10, %12, %4;\n madc.hi.cc.u32 %5, %10, %12, %5;\n"
            "madc.lo.cc.u32 %6, %11, %12, %6;\n madc.hi.u32 %7, %11, %12, %7;\n"
            : "+r"(e[0]), "+r"(e[1]), "+r"(e[2]), "+r"(e[3]), "+r"(e[4]), "+r"(e[5]), "+r"(e[6]), "+r"(e[7])
            : "r"(a[0]), "r"(a[2]), "r"(a[4]), "r"(a[6]), "r"(bi));
        asm volatile(
            "mad.lo.cc.u32 %0, %8, %12, %0;\n madc.hi.cc.u32 %1, %8, %12, %1;\n"
            "madc.lo.cc.u32 %2, %9, %12, %2;\n madc.hi.cc.u32 %3, %9, %12, %3;\n"
            "madc.lo.cc.u32 %4, %10, %12, %4;\n madc.hi.cc.u32 %5, %10, %12, %5;\n"
            "madc.lo.cc.u32 %6, %11, %12, %6;\n madc.hi.u32 %7, %11, %12, %7;\n"
            : "+r"(o[0]), "+r"(o[1]), "+r"(o[2]), "+r"(o[3]), "+r"(o[4]), "+r"(o[5]), "+r"(o[6]), "+r"(o[7])
            : "r"(a[1]), "r"(a[3]), "r"(a[5]), "r"(a[7]), "r"(bi));
    }
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s ^= e[i] ^ o[i];
    if (s == 0x1234567u) *sink = s;
}
// 7: same chains with immediate multiplicands (the modulus rows)
__global__ void k_chain_imm(uint32_t seed, unsigned long long* sink) {
    uint32_t e[8], o[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { e[i] = seed * (i + 1) + threadIdx.x; o[i] = e[i] * 3u; }
    for (int it = 0; it < ITERS; it++) {
        uint32_t bi = e[1] | 1u;
        asm volatile(
            "mad.lo.cc.u32 %0, %8, 0xf0000001, %0;\n madc.hi.cc.u32 %1, %8, 0xf0000001, %1;\n"
            "madc.lo.cc.u32 %2, %8, 0x79b97091, %2;\n madc.hi.cc.u32 %3, %8, 0x79b97091, %3;\n"
            "madc.lo.cc.u32 %4, %8, 0x8181585d, %4;\n madc.hi.cc.u32 %5, %8, 0x8181585d, %5;\n"
            "madc.lo.cc.u32 %6, %8, 0xe131a029, %6;\n madc.hi.u32 %7, %8, 0xe131a029, %7;\n"
            : "+r"(e[0]), "+r"(e[1]), "+r"(e[2]), "+r"(e[3]), "+r"(e[4]), "+r"(e[5]), "+r"(e[6]), "+r"(e[7])
            : "r"(bi));
        asm volatile(
            "mad.lo.cc.u32 %0, %8, 0x43e1f593, %0;\n madc.hi.cc.u32 %1, %8, 0x43e1f593, %1;\n"
            "madc.lo.cc.u32 %2, %8, 0x2833e848, %2;\n madc.hi.cc.u32 %3, %8, 0x2833e848, %3;\n"
            "madc.lo.cc.u32 %4, %8, 0xb85045b6, %4;\n madc.hi.cc.u32 %5, %8, 0xb85045b6, %5;\n"
            "madc.lo.cc.u32 %6, %8, 0x30644e72, %6;\n madc.hi.u32 %7, %8, 0x30644e72, %7;\n"
            : "+r"(o[0]), "+r"(o[1]), "+r"(o[2]), "+r"(o[3]), "+r"(o[4]), "+r"(o[5]), "+r"(o[6]), "+r"(o[7])
            : "r"(bi));
    }
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s ^= e[i] ^ o[i];
    if (s == 0x1234567u) *sink = s;
}
// 8: IADD3 only (ALU pipe) for reference
__global__ void k_iadd3(uint32_t seed, unsigned long long* sink) {
    uint32_t acc[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = seed * (i + 1) + threadIdx.x; b[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("add.u32 %0, %0, %1; add.u32 %0, %0, %2;" : "+r"(acc[i]) : "r"(acc[(i + 3) & 7]), "r"(b[i]));
    }
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s ^= acc[i];
    if (s == 0x1234567u) *sink = s;
}

template <class K>
static void run(const char* name, K kern, double instr_per_iter, int threads, int blocks_per_sm) {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    unsigned long long* sink; cudaMalloc(&sink, 8);
    int blocks = p.multiProcessorCount * blocks_per_sm;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    float best = 1e9f;
    for (int rep = 0; rep < 4; rep++) {
        cudaEventRecord(a);
        kern<<<blocks, threads>>>(12345u + rep, sink);
        cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b);
        if (rep && ms < best) best = ms;
    }
    double warp_instr = (double)blocks * threads / 32.0 * ITERS * instr_per_iter;
    int clk_khz; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    double cyc = best * 1e-3 * clk_khz * 1e3;   // at max clock
    printf("%-16s thr=%4d bps=%d  %.3f ms  %.3e thread-instr/s  %.2f cyc/warp-instr/SMSP (at %d MHz)\n", name, threads, blocks_per_sm,
           best, warp_instr * 32 / (best * 1e-3), cyc * p.multiProcessorCount * 4 / warp_instr, clk_khz / 1000);
    cudaFree(sink);
}

int main() {
    for (int bps : {2, 8}) {
        run("wide_indep", k_wide_indep, 8, 256, bps);
        run("wide_shared", k_wide_shared, 8, 256, bps);
        run("wide_imm", k_wide_imm, 8, 256, bps);
        run("lo32", k_lo32, 8, 256, bps);
        run("hi32", k_hi32, 8, 256, bps);
        run("chain_reg(16)", k_chain_reg, 16, 256, bps);
        run("chain_imm(16)", k_chain_imm, 16, 256, bps);
        run("iadd3(16)", k_iadd3, 16, 256, bps);
    }
    return 0;
}
