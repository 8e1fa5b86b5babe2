// msm_kernels.hpp -- the MSM / calibration kernels are compiled in msm_kernels.cu (their device code takes minutes to
// build); the orchestration unit (h2b200.cu) launches them through this table of __global__ entry points.
#pragma once
#include "msm.cuh"

namespace h2b {
struct MsmKernelTable {
    void (*count)(MsmGeom, const Fr*, uint32_t*, Fr*);
    void (*scatter)(MsmGeom, const Fr*, const Fr*, uint32_t*, uint32_t*);
    void (*scan_block_sums)(const uint32_t*, uint32_t, uint32_t*);
    void (*scan_single_block)(uint32_t*, uint32_t, uint32_t*);
    void (*scan_apply)(const uint32_t*, uint32_t, const uint32_t*, uint32_t*, uint32_t*);
    void (*precompute)(const Affine*, Affine*, uint32_t, uint32_t, uint32_t, uint32_t);
    void (*size_hist)(const uint32_t*, const uint32_t*, uint32_t, uint32_t*);
    void (*size_scatter)(const uint32_t*, const uint32_t*, uint32_t, uint32_t*, uint32_t*);
    void (*accumulate[4])(const Affine*, const uint32_t*, const uint32_t*, const uint32_t*, const uint32_t*, uint32_t, uint32_t,
                          Xyzz*, uint32_t);
    void (*long_list)(const uint32_t*, const uint32_t*, uint32_t, uint32_t, MsmLongInfo*, MsmLongTask*, uint32_t*);
    void (*accumulate_long)(const Affine*, const uint32_t*, const uint32_t*, const uint32_t*, const MsmLongTask*, const uint32_t*,
                            Xyzz*);
    void (*long_combine)(const MsmLongInfo*, const uint32_t*, const Xyzz*, Xyzz*, uint32_t);
    void (*reduce)(MsmGeom, const Xyzz*, uint32_t, uint32_t, Xyzz*, Xyzz*);
    void (*reduce_final)(uint32_t, uint32_t, const Xyzz*, const Xyzz*, Xyzz*);
    void (*reduce_small)(MsmGeom, const Xyzz*, Xyzz*);
    void (*reduce_lines)(const Xyzz*, uint32_t, uint32_t, uint32_t, uint32_t, Xyzz*);
    void (*reduce_masked)(const Xyzz*, uint32_t, uint32_t, uint32_t, uint32_t, Xyzz*);
    void (*peer_exchange)(const Xyzz*, uint32_t, MsmPeerTable, uint32_t, uint32_t, uint32_t, unsigned long long, unsigned long long,
                          unsigned long long, Xyzz*, uint32_t*);
    void (*imad_wide_peak)(uint32_t, uint32_t, unsigned long long*);
    void (*field_mul_peak)(uint32_t, Fq*);
    void (*synth_bases)(Affine*, uint64_t, uint64_t);
};
const MsmKernelTable& msm_kernel_table();
}  // namespace h2b
