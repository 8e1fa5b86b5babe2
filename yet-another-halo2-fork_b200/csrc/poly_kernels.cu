// poly_kernels.cu -- device code of poly.cuh, exported to h2b200.cu as a table of kernel entry points.
#include <cuda_runtime.h>

#include "poly_kernels.hpp"

namespace h2b {
const PolyKernelTable& poly_kernel_table() {
    static const PolyKernelTable t = {poly_prod_tile_reduce_kernel, poly_prod_scan_totals_kernel, poly_prod_tile_scan_kernel,
                                      poly_invert_totals_kernel,    poly_batch_invert_kernel,     poly_horner_tile_reduce_kernel, poly_horner_scan_totals_kernel,
                                      poly_horner_tile_scan_kernel, poly_horner_batch_tile_reduce_kernel, poly_horner_batch_totals_kernel, poly_mul_kernel};
    return t;
}
}  // namespace h2b
