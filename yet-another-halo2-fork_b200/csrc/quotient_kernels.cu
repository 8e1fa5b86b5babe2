// quotient_kernels.cu -- device code of quotient.cuh, exported to h2b200.cu as a table of kernel entry points.
#include <cuda_runtime.h>

#include "quotient_kernels.hpp"

namespace h2b {
const QuotientKernelTable& quotient_kernel_table() {
    static const QuotientKernelTable t = {quotient_graph_kernel, quotient_permutation_kernel, quotient_lookup_kernel,
                                          quotient_shuffle_kernel};
    return t;
}
}  // namespace h2b
