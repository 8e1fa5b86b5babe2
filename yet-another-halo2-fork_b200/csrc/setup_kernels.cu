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

// The vanishing argument's random polynomial (vanishing/prover.rs:57-81): stream s = ChaCha20Rng::from_seed(seeds[s]) fills
// out[offsets[s] .. offsets[s+1]) with consecutive Fr::random draws.  One draw = eight next_u64 = the sixteen words of one ChaCha20
// block (a fresh BlockRng is block aligned), so element j of a stream is from_u512(block j): one thread per element.
// rand_chacha layout: words 0-3 "expand 32-byte k", 4-11 key, 12-13 the 64-bit block counter, 14-15 the stream id (0).
#define H2B_QR(a, b, c, d)                                   \
    a += b; d ^= a; d = __funnelshift_l(d, d, 16);           \
    c += d; b ^= c; b = __funnelshift_l(b, b, 12);           \
    a += b; d ^= a; d = __funnelshift_l(d, d, 8);            \
    c += d; b ^= c; b = __funnelshift_l(b, b, 7);

__global__ void fr_random_chacha20_kernel(const uint32_t* __restrict__ seeds, const uint64_t* __restrict__ offsets,
                                          uint32_t n_streams, uint64_t n, Fr* __restrict__ out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t lo = 0, hi = n_streams;                    // the stream whose range holds i
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (offsets[mid] <= i) lo = mid; else hi = mid;
    }
    const uint64_t ctr = i - offsets[lo];
    uint32_t in[16];
    in[0] = 0x61707865u; in[1] = 0x3320646eu; in[2] = 0x79622d32u; in[3] = 0x6b206574u;
#pragma unroll
    for (int j = 0; j < 8; j++) in[4 + j] = seeds[lo * 8 + j];
    in[12] = (uint32_t)ctr; in[13] = (uint32_t)(ctr >> 32); in[14] = 0; in[15] = 0;
    uint32_t x0 = in[0], x1 = in[1], x2 = in[2], x3 = in[3], x4 = in[4], x5 = in[5], x6 = in[6], x7 = in[7], x8 = in[8],
             x9 = in[9], x10 = in[10], x11 = in[11], x12 = in[12], x13 = in[13], x14 = in[14], x15 = in[15];
    for (int round = 0; round < 10; round++) {
        H2B_QR(x0, x4, x8, x12) H2B_QR(x1, x5, x9, x13) H2B_QR(x2, x6, x10, x14) H2B_QR(x3, x7, x11, x15)
        H2B_QR(x0, x5, x10, x15) H2B_QR(x1, x6, x11, x12) H2B_QR(x2, x7, x8, x13) H2B_QR(x3, x4, x9, x14)
    }
    uint32_t w[16] = {x0 + in[0], x1 + in[1], x2 + in[2],  x3 + in[3],  x4 + in[4],  x5 + in[5],  x6 + in[6],  x7 + in[7],
                      x8 + in[8], x9 + in[9], x10 + in[10], x11 + in[11], x12 + in[12], x13 + in[13], x14 + in[14], x15 + in[15]};
    Fr l = reduce256(w), h = reduce256(w + 8);
    Fr r2 = Fr::r2();
    fe_store(out + i, fe_add(fe_mul(l, r2), fe_mul(fe_mul(h, r2), r2)));
}

}  // namespace

const SetupKernelTable& setup_kernel_table() {
    static const SetupKernelTable t = {g1_fixed_base_mul_kernel, fr_from_u512_kernel, fr_random_chacha20_kernel};
    return t;
}
}  // namespace h2b
