// lookup_kernels.hpp -- permute_expression_pair of the lookup argument (halo2_backend/src/plonk/lookup/prover.rs:398-494) on the
// device: the reference sorts the compressed input column, walks it against a BTreeMap of table multiplicities and fills the repeated
// rows with the left-over table values from the back.  Here: a radix sort of the canonical values (cub::DeviceRadixSort on one 64-bit
// limb per pass, only over the bits that differ between keys; the most significant varying limb alone when no two keys tie on it,
// otherwise least-significant-limb-first over every varying limb), dense ranks by a flag scan, a histogram of the table ranks, and
// one emit kernel.  Compiled in
// lookup_kernels.cu (the only translation unit that includes CUB).
#pragma once
#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

#include "field.cuh"

namespace h2b {
// bytes of device scratch lookup_permute_run needs for `usable` rows
size_t lookup_permute_scratch_bytes(uint32_t usable);
// Enqueues everything on `stream`, synchronising it once or twice on the way (the sort plan is made on the host from 64 bytes of
// key statistics).  d_status (one uint32 inside the scratch, returned through *d_status_out) ends up non-zero when
// an input value is missing from the table (bit 0) or the multiset sizes disagree (bit 1).  *launches counts the kernels enqueued.
cudaError_t lookup_permute_run(cudaStream_t stream, const Fr* d_input, const Fr* d_table, uint32_t usable, uint64_t n,
                               Fr* d_out_input, Fr* d_out_table, void* d_scratch, size_t scratch_bytes, uint32_t** d_status_out,
                               int* launches);
}  // namespace h2b
