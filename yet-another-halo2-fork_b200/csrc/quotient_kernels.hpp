// quotient_kernels.hpp -- entry points of the quotient-evaluation kernels (quotient.cuh), compiled in quotient_kernels.cu.
#pragma once
#include "quotient.cuh"

namespace h2b {
struct QuotientKernelTable {
    void (*graph)(const GraphArgs);
    void (*permutation)(const PermFoldArgs);
    void (*lookup)(const LookupFoldArgs);
    void (*shuffle)(const ShuffleFoldArgs);
};
const QuotientKernelTable& quotient_kernel_table();
}  // namespace h2b
