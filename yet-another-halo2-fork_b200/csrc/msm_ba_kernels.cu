// msm_ba_kernels.cu -- device code of the batched-affine bucket accumulation (msm_ba.cuh), exported as a table of entry points.
#include <cuda_runtime.h>

#define H2B_BA_KERNELS 1
#include "msm_ba_kernels.hpp"

namespace h2b {
const MsmBaKernelTable& msm_ba_kernel_table() {
    static const MsmBaKernelTable t = {{{msm_accumulate_ba_kernel<1, 4>, msm_accumulate_ba_kernel<1, 3>},
                                        {msm_accumulate_ba_kernel<2, 4>, msm_accumulate_ba_kernel<2, 3>}}};
    return t;
}
}  // namespace h2b
