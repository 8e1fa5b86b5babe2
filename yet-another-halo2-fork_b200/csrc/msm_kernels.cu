// msm_kernels.cu -- device code of the MSM pipeline (msm.cuh), the integer-pipe calibration kernels and the synthetic
// base generator, exported to h2b200.cu as a table of kernel entry points (msm_kernels.hpp).
#include <cuda_runtime.h>

#include "msm_kernels.hpp"

namespace h2b {
namespace {
// ------------------------------------------------------------------------------------------------
// integer-pipe calibration kernels
// ------------------------------------------------------------------------------------------------
// IMAD.WIDE carry chains exactly as the field multiplier issues them (two interleaved chains of four
// lo/hi pairs sharing one multiplier register; per-thread operands so nothing lands on the uniform datapath).
// Counts one "wide mad" per IMAD.WIDE.U32(.X): 8 per iteration.
__global__ void imad_wide_peak_kernel(uint32_t seed, uint32_t iters, unsigned long long* sink) {
    uint32_t e[8], o[8], a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) {
        e[i] = seed * (i + 1) + threadIdx.x;
        o[i] = e[i] * 3u + blockIdx.x;
        a[i] = seed + 77u * i + blockIdx.x + threadIdx.x * 2654435761u;
    }
    for (uint32_t it = 0; it < iters; it++) {
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
#pragma unroll
    for (int i = 0; i < 8; i++) s ^= e[i] ^ o[i];
    if (s == 0x1234567u) *sink = s;
}
// Two independent chains of dependent field multiplications per thread, all operands per-thread.
__global__ void field_mul_peak_kernel(uint32_t iters, Fq* sink) {
    Fq x = Fq::one(), y = Fq::r2();
#pragma unroll
    for (int i = 0; i < 7; i++) { x.l[i] ^= threadIdx.x * 2654435761u + i; y.l[i] ^= (threadIdx.x + 7u * blockIdx.x) * 40503u + i; }
    Fq a = x, b = y;
    for (uint32_t it = 0; it < iters; it++) {
        a = fe_mul(a, y);
        b = fe_mul(b, x);
        x.l[0] ^= b.l[3] & 1u;          // keep the operands loop-variant
    }
    Fq r = fe_add(a, b);
    if (r.l[0] == 0x12345u && r.l[7] == 0x54321u) *sink = r;
}

// Synthetic benchmark bases: out[i] = [t0 + i] G, G = (1, 2), affine.
__global__ void synth_bases_kernel(Affine* out, uint64_t n, uint64_t t0) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint64_t k = t0 + i;
    Fq gx = fe_from_u64<FqParams>(1), gy = fe_from_u64<FqParams>(2);
    Xyzz acc = Xyzz::identity();
    for (int bit = 63; bit >= 0; bit--) {
        acc = xyzz_double(acc);
        if ((k >> bit) & 1ull) xyzz_madd(acc, gx, gy, false);
    }
    Affine a = xyzz_to_affine(acc);
    fe_store(&out[i].x, a.x);
    fe_store(&out[i].y, a.y);
}

}  // namespace

const MsmKernelTable& msm_kernel_table() {
    static const MsmKernelTable t = {
        msm_count_kernel, msm_scatter_kernel, scan_block_sums_kernel, scan_single_block_kernel, scan_apply_kernel,
        msm_precompute_kernel, msm_size_hist_kernel, msm_size_scatter_kernel,
        {msm_accumulate_kernel<4, false>, msm_accumulate_kernel<4, true>, msm_accumulate_kernel<5, true>, msm_accumulate_kernel<6, true>},
        msm_long_list_kernel, msm_accumulate_long_kernel, msm_long_combine_kernel, msm_reduce_kernel, msm_reduce_final_kernel,
        msm_reduce_small_kernel, msm_reduce_lines_kernel, msm_reduce_masked_kernel, msm_peer_exchange_kernel, imad_wide_peak_kernel, field_mul_peak_kernel, synth_bases_kernel};
    return t;
}
}  // namespace h2b
