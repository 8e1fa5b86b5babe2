// noalloc.cu -- h2b_batch_invert_device_noalloc: batch inversion that never asks the driver for memory.
//
// Who this is for.  Every other entry point of the library works on an h2b_ctx, and a context allocates: h2b_init creates
// staging rings and pinned buffers, and each call grows the context's device scratch on demand (the first batch inversion of
// a new size ends up in cudaMalloc).  Some hosts cannot live with that:
//   * a prover service that takes (nearly) all of HBM into its own arena at start-up -- SRS, proving key, column pools -- has
//     nothing left for the driver to hand out, so a later cudaMalloc simply fails;
//   * a pipeline with other streams in flight does not want the device-wide stall a cudaMalloc / cudaFree may cause in the
//     middle of a proof;
//   * a component that owns no context (and must not create one) but holds a device buffer and a stream.
// For them the tile scratch of the inversion lives in the library image itself: a file-scope __device__ array, present once on
// every GPU that loads libh2b200.so, from load until the process exits.  It costs H2B_NOALLOC_SCRATCH_BYTES (2 MiB) per GPU
// whether or not the entry point is ever called, which is why the size limit is modest.
//
// The kernels are the ones h2b_batch_invert_device launches (poly.cuh: tile products, one block inverting the tile products
// with a single Fermat inversion, tiles inverting their elements); only where the tile scratch comes from differs, so the
// results are identical bit for bit.  Nothing in the scratch has to be zero when a call starts: the first kernel writes every
// tile product before the second reads any, and the second writes its prefix area before reading it.  The array keeps whatever
// the previous call left there.
#include <cuda_runtime.h>

#include <mutex>

#include "../../include/h2b200.h"
#define H2B_NO_KERNELS 1      // the kernels are compiled in poly_kernels.cu; this unit only launches them
#include "poly_kernels.hpp"

namespace {
using namespace h2b;

constexpr size_t kMaxElems = (size_t)H2B_NOALLOC_MAX_ELEMS;
constexpr size_t kMaxTiles = (kMaxElems + POLY_TILE - 1) / POLY_TILE;
constexpr size_t kScratchBytes = kMaxTiles * 2 * sizeof(Fr);      // tile products + the per-thread prefixes of the totals kernel
static_assert(kScratchBytes == (size_t)H2B_NOALLOC_SCRATCH_BYTES, "header and scratch array disagree");

// The scratch.  In HOST code this name is only a stand-in the runtime uses to look the array up; it is not a device address and
// must never be handed to a kernel as one.  The address each GPU gave its copy is fetched below with cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;                               // one call at a time enqueues its three kernels (they share the scratch)
void* g_scratch_addr[kMaxDevices] = {};        // per device: where that GPU holds g_noalloc_scratch (0 = not looked up yet)

int fail(const char** error_out, int code, const char* msg) {
    if (error_out) *error_out = msg;
    return code;
}
}  // namespace

extern "C" int h2b_batch_invert_device_noalloc(void* d_a, size_t n, void* cuda_stream, const char** error_out) {
    if (error_out) *error_out = "";
    if (n == 0) return H2B_OK;
    if (!d_a) return fail(error_out, H2B_E_INVALID, "batch_invert_noalloc: null argument");
    if (n > kMaxElems)
        return fail(error_out, H2B_E_INVALID,
                    "batch_invert_noalloc: n exceeds H2B_NOALLOC_MAX_ELEMS (2^26 elements), the size the built-in scratch "
                    "was laid out for; use h2b_batch_invert_device, or split the input");
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) { cudaGetLastError(); return fail(error_out, H2B_E_NODEVICE, "batch_invert_noalloc: no current CUDA device"); }
    if (dev < 0 || dev >= kMaxDevices) return fail(error_out, H2B_E_INVALID, "batch_invert_noalloc: device ordinal beyond 63");

    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_addr[dev]) {
        // Set-up, once per device, with that device current: ask where THIS GPU keeps its copy of the array.
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(error_out, H2B_E_CUDA, "batch_invert_noalloc: cudaGetSymbolAddress failed (is this an sm_100 device?)");
        }
        g_scratch_addr[dev] = p;
    }
    const PolyKernelTable& PK = poly_kernel_table();
    cudaStream_t st = (cudaStream_t)cuda_stream;
    Fr* a = (Fr*)d_a;
    Fr* tt = (Fr*)g_scratch_addr[dev];
    const uint32_t tiles = (uint32_t)((n + POLY_TILE - 1) / POLY_TILE);
    PK.prod_tile_reduce<<<tiles, POLY_T, 0, st>>>((const Fr*)a, (uint64_t)n, tt, 1u);
    PK.invert_totals<<<1, POLY_T, 0, st>>>(tt, tt + tiles, tiles);
    PK.batch_invert<<<tiles, POLY_T, 0, st>>>(a, (uint64_t)n, (const Fr*)tt);
    if (cudaGetLastError() != cudaSuccess) return fail(error_out, H2B_E_CUDA, "batch_invert_noalloc: kernel launch failed");
    return H2B_OK;
}
