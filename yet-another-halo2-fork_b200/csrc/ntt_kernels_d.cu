// ntt_kernels_d.cu -- instantiates ntt_pass_kernel for log_s in [9] (see ntt_kernels.hpp).
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
void h2b_ntt_register_d(ntt_kernel_t tab[10][2], size_t smem[10]) {
    reg<9>(tab, smem);
}
}  // namespace h2b
