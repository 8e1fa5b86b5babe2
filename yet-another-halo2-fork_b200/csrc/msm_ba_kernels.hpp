// msm_ba_kernels.hpp -- entry points of the batched-affine bucket accumulation (msm_ba.cuh), compiled in msm_ba_kernels.cu.
#pragma once
#include "msm_ba.cuh"

namespace h2b {
struct MsmBaKernelTable {
    // [rounds - 1][0: 4 resident blocks per SM (128 registers), 1: 3 blocks (168 registers)]
    void (*accumulate[2][2])(const Affine*, const uint32_t*, const uint32_t*, const uint32_t*, const uint32_t*, uint32_t, uint32_t,
                             Xyzz*, uint32_t, Fq*, Fq*, Fq*, uint32_t);
};
const MsmBaKernelTable& msm_ba_kernel_table();
}  // namespace h2b
