// poly_kernels.hpp -- entry points of the scan-shaped polynomial kernels (poly.cuh), compiled in poly_kernels.cu.
#pragma once
#include "poly.cuh"

namespace h2b {
struct PolyKernelTable {
    void (*prod_tile_reduce)(const Fr*, uint64_t, Fr*, uint32_t);
    void (*prod_scan_totals)(Fr*, uint32_t, Fr);
    void (*prod_tile_scan)(const Fr*, uint64_t, const Fr*, Fr*, uint64_t);
    void (*invert_totals)(Fr*, Fr*, uint32_t);
    void (*batch_invert)(Fr*, uint64_t, const Fr*);
    void (*horner_tile_reduce)(const Fr*, uint64_t, uint64_t, Fr, Fr, Fr*);
    void (*horner_scan_totals)(Fr*, uint32_t, Fr, Fr*);
    void (*horner_tile_scan)(const Fr*, uint64_t, uint64_t, Fr, Fr, const Fr*, Fr*);
    void (*horner_batch_tile_reduce)(const HornerJob*, uint64_t, uint64_t, uint32_t, Fr*);
    void (*horner_batch_totals)(const HornerJob*, const Fr*, uint32_t, Fr*);
    void (*mul)(const Fr*, const Fr*, Fr, uint32_t, Fr*, uint64_t);
};
const PolyKernelTable& poly_kernel_table();
}  // namespace h2b
