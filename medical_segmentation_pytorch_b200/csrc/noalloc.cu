// msx_bn_act_bwd_static: BatchNorm(+ReLU) backward for callers that must not allocate.
//
// Some users of this library cannot hand us a workspace: code that runs inside another framework's
// memory plan where every device byte is already spoken for, real-time or safety-reviewed deployments
// in which cudaMalloc after start-up is forbidden because it may block, fragment or fail, and plug-in
// sandboxes that only ever see the input and output buffers they were given.  For them the partial
// sums of the BatchNorm backward reduction live in a buffer that is part of the library image itself.
//
// The price: the buffer is a fixed size (so the channel count is capped), every GPU that loads
// libmsx.so carries it until the process exits, and all calls share it.
#include <mutex>

#include "common.cuh"

namespace msx {
namespace {

// Room for MSX_BN_STATIC_MAX_CHANNELS channels: 592 per-CTA partial rows x 2 sums x C doubles, plus the two
// per-channel fp32 coefficient vectors the apply pass reads, plus slack (the layout bn_bwd_launch uses).
constexpr int kStaticMaxChannels = MSX_BN_STATIC_MAX_CHANNELS;
constexpr size_t kStaticBytes = (size_t)kMaxPartialBlocks * 2 * kStaticMaxChannels * sizeof(double) +
                                2 * (size_t)kStaticMaxChannels * sizeof(float) + 256;

// Device-resident scratch.  NOTE for maintainers: in host code this name is only a handle the runtime
// knows the symbol by - it is NOT a device pointer and must never be passed to a kernel as one (the
// kernel would fault, or on systems where the GPU can reach host memory quietly scribble over the wrong
// bytes).  Host code obtains the real address with cudaGetSymbolAddress, below.
static __device__ __align__(256) unsigned char g_bn_bwd_scratch[kStaticBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
void* g_scratch_ptr[kMaxDevices] = {nullptr};

// Looks the scratch address up once per device.  The device must be current: a __device__ array exists
// separately on every GPU and cudaGetSymbolAddress answers for the current one.
int scratch_for_current_device(void** out) {
    int dev = -1;
    MSX_CHECK_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return set_error(MSX_ERR_UNSUPPORTED, "bn_act_bwd_static: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_ptr[dev]) {
        void* p = nullptr;
        MSX_CHECK_CUDA(cudaGetSymbolAddress(&p, g_bn_bwd_scratch));
        if (!p) return set_error(MSX_ERR_CUDA, "bn_act_bwd_static: cudaGetSymbolAddress returned a null address");
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return MSX_OK;
}

}  // namespace
}  // namespace msx

using namespace msx;

extern "C" int msx_bn_act_bwd_static(int dtype, const void* s, int32_t lds, const void* dz, int32_t lddz, int64_t pixels,
                                     int32_t channels, const float* gamma, const float* beta, const float* save_mean,
                                     const float* save_invstd, int relu, void* ds, int32_t ldds, float* dgamma,
                                     float* dbeta, void* stream) {
    MSX_REQUIRE(s && dz && ds && gamma && beta && save_mean && save_invstd, "bn_act_bwd_static: null pointer");
    MSX_REQUIRE(channels >= 1 && pixels >= 1, "bn_act_bwd_static: empty input");
    if (channels > kStaticMaxChannels)
        return set_error(MSX_ERR_UNSUPPORTED,
                         "bn_act_bwd_static: %d channels exceed the built-in scratch (MSX_BN_STATIC_MAX_CHANNELS = %d); "
                         "use msx_bn_act_bwd with a caller-provided workspace",
                         channels, kStaticMaxChannels);
    void* scratch = nullptr;
    MSX_TRY(scratch_for_current_device(&scratch));
    // Nothing in the scratch needs clearing between calls: the reduction pass writes every partial row it
    // later reads (rows [0, grid) x 2 x round_up(C, 8)) and the finalize pass writes both coefficient vectors
    // before the apply pass reads them, so stale contents of an earlier call are never consumed.
    return bn_bwd_launch(dtype, s, lds, dz, lddz, pixels, channels, gamma, beta, save_mean, save_invstd, relu, ds, ldds,
                         dgamma, dbeta, scratch, kStaticBytes, as_stream(stream));
}
