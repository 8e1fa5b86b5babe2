// FP64-pipe microbenchmarks for sm_100a: can DFMA (52x52-bit exact products, Emmart's trick) run beside IMAD.WIDE?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_micro fp64_micro.cu ; run on the GPU box.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 2048

__global__ void k_dfma(uint32_t seed, unsigned long long* sink) {
    double acc[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = (double)(seed * (i + 1) + threadIdx.x); b[i] = 1.0 + 1e-9 * (double)(seed + 77u * i + threadIdx.x); }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) acc[i] = __fma_rz(acc[i], b[i], b[(i + 1) & 7]);
    }
    double s = 0;
    for (int i = 0; i < 8; i++) s += acc[i];
    if (s == 0.1234567) *sink = (unsigned long long)s;
}
__global__ void k_dadd(uint32_t seed, unsigned long long* sink) {
    double acc[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = (double)(seed * (i + 1) + threadIdx.x); b[i] = 1.0 + 1e-9 * (double)(seed + 77u * i + threadIdx.x); }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) acc[i] = __dadd_rz(acc[i], b[i]);
    }
    double s = 0;
    for (int i = 0; i < 8; i++) s += acc[i];
    if (s == 0.1234567) *sink = (unsigned long long)s;
}
// IMAD.WIDE carry chain (8 mad = 4 IMAD.WIDE) + N DFMA per iteration, independent: do they overlap?
template <int NF>
__global__ void k_mix(uint32_t seed, unsigned long long* sink) {
    uint32_t e[8], a[8];
    double acc[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        e[i] = seed * (i + 1) + threadIdx.x; a[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u;
        acc[i] = (double)(seed * (i + 1) + threadIdx.x); b[i] = 1.0 + 1e-9 * (double)(seed + 77u * i + threadIdx.x);
    }
    for (int it = 0; it < ITERS; it++) {
        uint32_t bi = e[1] | 1u;
        asm volatile(
            "mad.lo.cc.u32 %0, %8, %12, %0;\n madc.hi.cc.u32 %1, %8, %12, %1;\n"
            "madc.lo.cc.u32 %2, %9, %12, %2;\n madc.hi.cc.u32 %3, %9, %12, %3;\n"
            "madc.lo.cc.u32 %4, %10, %12, %4;\n madc.hi.cc.u32 %5, %10, %12, %5;\n"
            "madc.lo.cc.u32 %6, %11, %12, %6;\n madc.hi.u32 %7, %11, %12, %7;\n"
            : "+r"(e[0]), "+r"(e[1]), "+r"(e[2]), "+r"(e[3]), "+r"(e[4]), "+r"(e[5]), "+r"(e[6]), "+r"(e[7])
            : "r"(a[0]), "r"(a[2]), "r"(a[4]), "r"(a[6]), "r"(bi));
#pragma unroll
        for (int i = 0; i < NF; i++) acc[i & 7] = __fma_rz(acc[i & 7], b[i & 7], b[(i + 1) & 7]);
    }
    uint32_t s = 0; double d = 0;
    for (int i = 0; i < 8; i++) { s ^= e[i]; d += acc[i]; }
    if (s == 0x1234567u && d == 0.5) *sink = s;
}
// IMAD.WIDE chain + N IADD3-ish (ALU pipe) per iteration
template <int NA>
__global__ void k_mix_alu(uint32_t seed, unsigned long long* sink) {
    uint32_t e[8], a[8], x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { e[i] = seed * (i + 1) + threadIdx.x; a[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u; x[i] = a[i] ^ 0x5555u; }
    for (int it = 0; it < ITERS; it++) {
        uint32_t bi = e[1] | 1u;
        asm volatile(
            "mad.lo.cc.u32 %0, %8, %12, %0;\n madc.hi.cc.u32 %1, %8, %12, %1;\n"
            "madc.lo.cc.u32 %2, %9, %12, %2;\n madc.hi.cc.u32 %3, %9, %12, %3;\n"
            "madc.lo.cc.u32 %4, %10, %12, %4;\n madc.hi.cc.u32 %5, %10, %12, %5;\n"
            "madc.lo.cc.u32 %6, %11, %12, %6;\n madc.hi.u32 %7, %11, %12, %7;\n"
            : "+r"(e[0]), "+r"(e[1]), "+r"(e[2]), "+r"(e[3]), "+r"(e[4]), "+r"(e[5]), "+r"(e[6]), "+r"(e[7])
            : "r"(a[0]), "r"(a[2]), "r"(a[4]), "r"(a[6]), "r"(bi));
#pragma unroll
        for (int i = 0; i < NA; i++) asm volatile("add.u32 %0, %0, %1; add.u32 %0, %0, %2;" : "+r"(x[i & 7]) : "r"(x[(i + 3) & 7]), "r"(a[i & 7]));
    }
    uint32_t s = 0;
    for (int i = 0; i < 8; i++) s ^= e[i] ^ x[i];
    if (s == 0x1234567u) *sink = s;
}
// 64-bit integer adds (what accumulating DFMA mantissas costs)
__global__ void k_add64(uint32_t seed, unsigned long long* sink) {
    unsigned long long acc[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = seed * (i + 1) + threadIdx.x; b[i] = ((unsigned long long)seed << 20) + 77u * i + threadIdx.x * 2654435761ull; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("add.u64 %0, %0, %1;" : "+l"(acc[i]) : "l"(b[(i + it) & 7] ^ acc[(i + 3) & 7]));
    }
    unsigned long long s = 0;
    for (int i = 0; i < 8; i++) s ^= acc[i];
    if (s == 0x1234567ull) *sink = s;
}
// funnel shifts (binary-GCD inversion would live on these)
__global__ void k_shf(uint32_t seed, unsigned long long* sink) {
    uint32_t acc[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { acc[i] = seed * (i + 1) + threadIdx.x; b[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u; }
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) acc[i] = __funnelshift_r(acc[i], acc[(i + 1) & 7], 1) ^ b[i];
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
    double warp_iters = (double)blocks * threads / 32.0 * ITERS;
    int clk_khz; cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0);
    double cyc = best * 1e-3 * clk_khz * 1e3;
    double cyc_per_iter = cyc * p.multiProcessorCount * 4 / warp_iters;
    printf("%-22s thr=%4d bps=%d  %.3f ms  %.2f cyc/warp-iteration/SMSP  (%.2f per counted instr, %g instr/iter)\n", name, threads, blocks_per_sm,
           best, cyc_per_iter, cyc_per_iter / instr_per_iter, instr_per_iter);
    cudaFree(sink);
}

int main() {
    for (int bps : {2, 8}) {
        run("dfma x8", k_dfma, 8, 256, bps);
        run("dadd x8", k_dadd, 8, 256, bps);
        run("imadwide x4 + dfma x0", k_mix<0>, 4, 256, bps);
        run("imadwide x4 + dfma x4", k_mix<4>, 4, 256, bps);
        run("imadwide x4 + dfma x8", k_mix<8>, 4, 256, bps);
        run("imadwide x4 + dfma x16", k_mix<16>, 4, 256, bps);
        run("imadwide x4 + iadd3 x8", k_mix_alu<4>, 4, 256, bps);
        run("imadwide x4 + iadd3 x16", k_mix_alu<8>, 4, 256, bps);
        run("imadwide x4 + iadd3 x24", k_mix_alu<12>, 4, 256, bps);
        run("add.u64 x8", k_add64, 8, 256, bps);
        run("shf x8 (+lop)", k_shf, 8, 256, bps);
    }
    return 0;
}
