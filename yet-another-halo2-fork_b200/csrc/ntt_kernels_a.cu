// ntt_kernels_a.cu -- instantiates ntt_pass_kernel for log_s in [1, 2, 3, 4, 5] (see ntt_kernels.hpp).
#include "ntt_kernels.hpp"

namespace h2b {
namespace {
template <int LOG_S>
void reg(ntt_kernel_t tab[10][2], size_t smem[10]) {
    tab[LOG_S][0] = ntt_pass_kernel<LOG_S, false>;
    tab[LOG_S][1] = ntt_pass_kernel<LOG_S, true>;
    smem[LOG_S] = NttSmem<LOG_S>::bytes;
}
}  // namespace
void h2b_ntt_register_a(ntt_kernel_t tab[10][2], size_t smem[10]) {
    reg<1>(tab, smem);
    reg<2>(tab, smem);
    reg<3>(tab, smem);
    reg<4>(tab, smem);
    reg<5>(tab, smem);
}
}  // namespace h2b
