// ntt_kernels.hpp -- the NTT pass kernels are instantiated in four translation units (ntt_kernels_{a,b,c,d}.cu)
// so that the library builds in parallel; each registers its instantiations into the dispatch table.
#pragma once
#include <cstddef>
#include "ntt.cuh"

namespace h2b {
typedef void (*ntt_kernel_t)(const NttPass, const Fr*, Fr*, uint64_t, uint64_t);
void h2b_ntt_register_a(ntt_kernel_t tab[10][2], size_t smem[10]);   // log_s 1..5
void h2b_ntt_register_b(ntt_kernel_t tab[10][2], size_t smem[10]);   // log_s 6, 7
void h2b_ntt_register_c(ntt_kernel_t tab[10][2], size_t smem[10]);   // log_s 8
void h2b_ntt_register_d(ntt_kernel_t tab[10][2], size_t smem[10]);   // log_s 9
}  // namespace h2b
