// msm_bin_kernels.cu -- device code of the two-level digit binning (msm_bin.cuh), exported as a table of entry points.
#include <cuda_runtime.h>

#define H2B_NO_KERNELS 1       // the other MSM kernels are compiled in msm_kernels.cu
#define H2B_BIN_KERNELS 1
#include "msm_bin.cuh"

namespace h2b {
namespace {
int prepare_bin_kernels() {
    cudaError_t e = cudaFuncSetAttribute(msm_bin_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MSM_BIN_SMEM_BUDGET);
    if (e != cudaSuccess) return (int)e;
    return (int)cudaFuncSetAttribute(msm_partition_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)msm_part_smem());
}
}  // namespace
const MsmBinKernelTable& msm_bin_kernel_table() {
    static const MsmBinKernelTable t = {msm_digits_kernel, msm_partition_kernel, msm_bin_kernel, prepare_bin_kernels};
    return t;
}
}  // namespace h2b
