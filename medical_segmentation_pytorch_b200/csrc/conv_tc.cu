// tcgen05 / TMEM / TMA implicit-GEMM convolution engine (bf16).  Placeholder until the kernel lands.
#include "common.cuh"
namespace msx {
bool conv_tc_supported(const msx_conv_desc*, bool) { return false; }
int conv_gather_tc(const msx_conv_desc*, bool, const void*, const void*, const float*, void*, cudaStream_t) {
    return set_error(MSX_ERR_UNSUPPORTED, "tcgen05 conv engine not built");
}
}  // namespace msx
