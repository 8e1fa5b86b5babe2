// setup_kernels.cu -- see setup_kernels.hpp.
#include <cuda_runtime.h>

#include "setup_kernels.hpp"

namespace h2b {
namespace {

// out[i] = [k_i] G, G = (1, 2), k_i given in Montgomery form.  One thread per scalar: MSB-first double-and-add in XYZZ,
// one inversion at the end.  ~3 300 field multiplications per point; a 2^21-point SRS costs ~0.1 s, once.
__global__ void g1_fixed_base_mul_kernel(const Fr* __restrict__ scalars, uint64_t n, Affine* __restrict__ out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Fr k = fe_from_mont(fe_load(scalars + i));
    Fq gx = fe_from_u64<FqParams>(1), gy = fe_from_u64<FqParams>(2);
    Xyzz acc = Xyzz::identity();
    for (int bit = 253; bit >= 0; bit--) {
        acc = xyzz_double(acc);
        if ((k.l[bit >> 5] >> (bit & 31)) & 1u) xyzz_madd(acc, gx, gy, false);
    }
    Affine a = xyzz_to_affine(acc);
    fe_store(&out[i].x, a.x);
    fe_store(&out[i].y, a.y);
}

// in: n x 16 little-endian 32-bit words (a 512-bit integer v = lo + 2^256 hi); out = v mod r in Montgomery form.
// lo, hi < 2^256 < 6 r are first brought below r by conditional subtraction (the multiplier expects reduced operands);
// then M(lo) = lo * R2 / R, M(hi * 2^256) = (hi * R2 / R) * R2 / R.
__device__ __forceinline__ Fr reduce256(const uint32_t* w) {
    uint32_t v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) v[j] = w[j];
    for (int it = 0; it < 6; it++) {
        uint32_t t[8];
        uint64_t borrow = 0;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            uint64_t d = (uint64_t)v[j] - FrParams::p(j) - borrow;
            t[j] = (uint32_t)d;
            borrow = (d >> 63) & 1u;
        }
        if (borrow) break;
#pragma unroll
        for (int j = 0; j < 8; j++) v[j] = t[j];
    }
    Fr r;
#pragma unroll
    for (int j = 0; j < 8; j++) r.l[j] = v[j];
    return r;
}

__global__ void fr_from_u512_kernel(const uint32_t* __restrict__ in, uint64_t n, Fr* __restrict__ out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t* w = in + i * 16;
    Fr lo = reduce256(w), hi = reduce256(w + 8);
    Fr r2 = Fr::r2();
    Fr res = fe_add(fe_mul(lo, r2), fe_mul(fe_mul(hi, r2), r2));
    fe_store(out + i, res);
}

}  // namespace

const SetupKernelTable& setup_kernel_table() {
    static const SetupKernelTable t = {g1_fixed_base_mul_kernel, fr_from_u512_kernel};
    return t;
}
}  // namespace h2b
