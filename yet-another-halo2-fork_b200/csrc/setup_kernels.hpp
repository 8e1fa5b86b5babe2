// setup_kernels.hpp -- kernels of the set-up path (not the proving hot path): the fixed-base multiplications ParamsKZG::setup
// performs (poly/kzg/commitment.rs:64-131: g[i] = [s^i] G, g_lagrange[i] = [L_i(s)] G) and the wide reduction behind
// Fr::random / Fr::from_uniform_bytes (512-bit little-endian integer mod r), compiled in setup_kernels.cu.
#pragma once
#include "curve.cuh"

namespace h2b {
struct SetupKernelTable {
    void (*fixed_base_mul)(const Fr*, uint64_t, Affine*);
    void (*fr_from_u512)(const uint32_t*, uint64_t, Fr*);
    void (*fr_random_chacha20)(const uint32_t*, const uint64_t*, uint32_t, uint64_t, Fr*);
};
const SetupKernelTable& setup_kernel_table();
}  // namespace h2b
