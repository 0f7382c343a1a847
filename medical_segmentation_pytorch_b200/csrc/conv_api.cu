// extern "C" convolution entry points: engine selection + the ConvBNAct composites.
#include <stdlib.h>

#include "common.cuh"

namespace msx {
struct GatherGeom;
int validate_desc(const msx_conv_desc* d);
size_t conv_packed_elems(const msx_conv_desc* d);
int conv_pack_launch(const msx_conv_desc* d, int mode, const float* w, void* packed, cudaStream_t st);
size_t conv_pack_plan_bytes(const msx_pack_item* items, int count);
int conv_pack_plan_build(const msx_pack_item* items, int count, void* table, size_t table_bytes, int* chunks_out);
int conv_pack_plan_run(int dtype, const void* table, int count, int chunks, double bytes, cudaStream_t st);
int conv_gather_simt(const msx_conv_desc* d, bool fwd_pass, const void* src, const void* packed, const ConvEpilogue& ep,
                     void* dst, cudaStream_t st);
size_t conv_wgrad_ws_bytes(const msx_conv_desc* d);
int conv_wgrad_simt(const msx_conv_desc* d, const void* x, const void* dy, float* dw, void* ws, size_t ws_bytes,
                    cudaStream_t st);
// tcgen05 engine (conv_tc.cu)
bool conv_tc_supported(const msx_conv_desc* d, bool fwd_pass);
bool conv_tc_gather_supported(const msx_conv_desc* d, bool fwd_pass);
int conv_gather_tc(const msx_conv_desc* d, bool fwd_pass, const void* src, const void* packed, const ConvEpilogue& ep,
                   void* dst, cudaStream_t st, bool* stats_fused);

size_t conv_wgrad_tc_ws_bytes(const msx_conv_desc* d);
int conv_wgrad_tc(const msx_conv_desc* d, const void* x, const void* dy, float* dw, void* ws, size_t ws_bytes, cudaStream_t st);

static bool want_tc(const msx_conv_desc* d) {
    return d->engine == MSX_ENGINE_TCGEN05 || (d->engine == MSX_ENGINE_AUTO && d->dtype == MSX_BF16);
}

static int conv_wgrad(const msx_conv_desc* d, const void* x, const void* dy, float* dw, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (want_tc(d) && conv_tc_supported(d, true)) return conv_wgrad_tc(d, x, dy, dw, ws, ws_bytes, st);
    return conv_wgrad_simt(d, x, dy, dw, ws, ws_bytes, st);
}

static int conv_gather(const msx_conv_desc* d, bool fwd_pass, const void* src, const void* packed, const ConvEpilogue& ep,
                       void* dst, cudaStream_t st, bool* stats_fused = nullptr) {
    if (stats_fused) *stats_fused = false;
    const bool want_tc = d->engine == MSX_ENGINE_TCGEN05 || (d->engine == MSX_ENGINE_AUTO && d->dtype == MSX_BF16);
    if (want_tc && conv_tc_gather_supported(d, fwd_pass)) return conv_gather_tc(d, fwd_pass, src, packed, ep, dst, st, stats_fused);
    if (d->engine == MSX_ENGINE_TCGEN05)
        return set_error(MSX_ERR_UNSUPPORTED, "conv: tcgen05 engine requested for a shape/dtype it does not support");
    return conv_gather_simt(d, fwd_pass, src, packed, ep, dst, st);
}
static int conv_gather(const msx_conv_desc* d, bool fwd_pass, const void* src, const void* packed, const float* bias,
                       void* dst, cudaStream_t st) {
    return conv_gather(d, fwd_pass, src, packed, conv_epilogue(bias), dst, st);
}

// MSX_FUSED_STATS=0: statistics by the stand-alone pass everywhere (measurement / bisecting switch, read once)
static bool fused_stats_enabled() {
    static const bool on = [] { const char* e = getenv("MSX_FUSED_STATS"); return !(e && atoi(e) == 0); }();
    return on;
}

static size_t ws_conv_bytes(const msx_conv_desc* d) {
    size_t a = conv_wgrad_ws_bytes(d);
    if (d->dtype == MSX_BF16 && conv_tc_supported(d, true)) {
        const size_t t = conv_wgrad_tc_ws_bytes(d);
        if (t > a) a = t;
    }
    const int cmax = d->cin > d->cout ? d->cin : d->cout;
    const size_t b = bn_partials_bytes(cmax);
    return (a > b ? a : b) + 256;
}
}  // namespace msx

using namespace msx;

extern "C" {

size_t msx_conv_packed_bytes(const msx_conv_desc* d) {
    if (!d) return 0;
    return conv_packed_elems(d) * (d->dtype == MSX_BF16 ? 2 : 4);
}
size_t msx_ws_bytes_conv(const msx_conv_desc* d) { return d ? ws_conv_bytes(d) : 0; }

int msx_conv_pack_weights(const msx_conv_desc* d, int mode, const float* w_master, void* packed, void* stream) {
    MSX_TRY(validate_desc(d));
    MSX_REQUIRE(w_master && packed, "pack_weights: null pointer");
    MSX_REQUIRE(mode == MSX_PACK_FPROP || mode == MSX_PACK_DGRAD, "pack_weights: bad mode %d", mode);
    return conv_pack_launch(d, mode, w_master, packed, as_stream(stream));
}

size_t msx_conv_pack_plan_bytes(const msx_pack_item* host_items, int32_t count) {
    if (!host_items || count <= 0) return 0;
    return conv_pack_plan_bytes(host_items, count);
}
int msx_conv_pack_plan_build(const msx_pack_item* host_items, int32_t count, void* table, size_t table_bytes, int32_t* chunks_out) {
    MSX_REQUIRE(host_items && count > 0 && table && chunks_out, "pack_plan_build: bad arguments");
    int chunks = 0;
    MSX_TRY(conv_pack_plan_build(host_items, count, table, table_bytes, &chunks));
    *chunks_out = chunks;
    return MSX_OK;
}
int msx_conv_pack_plan_run(int dtype, const void* table, int32_t count, int32_t chunks, void* stream) {
    MSX_REQUIRE(table && count > 0 && chunks > 0, "pack_plan_run: bad arguments");
    MSX_REQUIRE(dtype == MSX_F32 || dtype == MSX_BF16, "pack_plan_run: bad dtype");
    // algorithmic bytes: every chunk reads <= 1024 pairs x taps floats and writes as many elements (upper bound)
    return conv_pack_plan_run(dtype, table, count, chunks, 0.0, as_stream(stream));
}

int msx_conv2d_fwd(const msx_conv_desc* d, const void* x, const float* w_master, const float* bias, void* w_packed,
                   void* y, void* stream) {
    MSX_TRY(validate_desc(d));
    MSX_REQUIRE(x && w_packed && y, "conv2d_fwd: null pointer");
    cudaStream_t st = as_stream(stream);
    if (w_master) MSX_TRY(conv_pack_launch(d, MSX_PACK_FPROP, w_master, w_packed, st));   // NULL: w_packed is pre-packed
    return conv_gather(d, true, x, w_packed, bias, y, st);
}

int msx_conv2d_bwd(const msx_conv_desc* d, const void* x, const float* w_master, const void* dy, void* w_packed,
                   void* dx, float* dw, float* dbias, void* ws, size_t ws_bytes, void* stream) {
    MSX_TRY(validate_desc(d));
    MSX_REQUIRE(x && dy && w_packed, "conv2d_bwd: null pointer");
    cudaStream_t st = as_stream(stream);
    if (dx) {
        if (w_master) MSX_TRY(conv_pack_launch(d, MSX_PACK_DGRAD, w_master, w_packed, st));
        MSX_TRY(conv_gather(d, false, dy, w_packed, nullptr, dx, st));
    }
    if (dw) MSX_TRY(conv_wgrad(d, x, dy, dw, ws, ws_bytes, st));
    if (dbias)
        MSX_TRY(channel_sum_launch(d->dtype, dy, d->ldy, (int64_t)d->n * d->oh * d->ow, d->cout, dbias, ws, ws_bytes, st));
    return MSX_OK;
}

int msx_conv_bn_act_fwd(const msx_conv_desc* d, const void* x, const float* w_master, const float* bias,
                        const float* gamma, const float* beta, float* running_mean, float* running_var,
                        int64_t* num_batches_tracked, float eps, float momentum, int training, int relu,
                        void* w_packed, void* y_raw, void* z, int32_t ldz, float* save_mean, float* save_invstd,
                        void* ws, size_t ws_bytes, void* stream) {
    MSX_TRY(validate_desc(d));
    MSX_REQUIRE(x && gamma && beta && w_packed && z && save_mean && save_invstd, "conv_bn_act_fwd: null pointer");
    MSX_REQUIRE(training || (running_mean && running_var), "conv_bn_act_fwd: eval mode needs running statistics");
    MSX_REQUIRE(y_raw || !training, "conv_bn_act_fwd: y_raw may only be NULL in eval mode");
    if (ws_bytes < bn_partials_bytes(d->cout)) return set_error(MSX_ERR_WORKSPACE, "conv_bn_act_fwd: workspace %zu < %zu", ws_bytes, bn_partials_bytes(d->cout));
    cudaStream_t st = as_stream(stream);
    const int64_t pixels = (int64_t)d->n * d->oh * d->ow;
    if (w_master) MSX_TRY(conv_pack_launch(d, MSX_PACK_FPROP, w_master, w_packed, st));
    if (!y_raw) {
        // inference (no backward will follow): eval-mode BN is an affine per channel, applied with the ReLU in the
        // conv epilogue straight from the fp32 accumulators - no raw conv output, no BN pass
        float* scale = static_cast<float*>(ws);
        float* shift = scale + round_up(d->cout, 8);
        MSX_TRY(bn_eval_affine_launch(d->cout, running_mean, running_var, eps, gamma, beta, bias, save_mean, save_invstd, scale, shift, st));
        msx_conv_desc dz = *d;
        dz.ldy = ldz;
        return conv_gather(&dz, true, x, w_packed, conv_epilogue(nullptr, scale, shift, relu), z, st);
    }
    if (training) {
        // batch statistics come out of the convolution's own epilogue (tcgen05 kernels: per-channel sums of what is stored,
        // finalized by the launch's last CTA); engines / geometries without that run one statistics pass over y_raw
        ConvEpilogue ep = conv_epilogue(bias);
        MSX_TRY(bn_finalize_args(&ep.bnf, static_cast<double*>(ws), d->cout, pixels, eps, momentum, save_mean, save_invstd, running_mean,
                                 running_var, num_batches_tracked, nullptr));
        ep.stats = fused_stats_enabled() ? 1 : 0;
        bool fused = false;
        MSX_TRY(conv_gather(d, true, x, w_packed, ep, y_raw, st, &fused));
        if (!fused) MSX_TRY(bn_stats_launch(d->dtype, y_raw, d->ldy, pixels, d->cout, static_cast<double*>(ws), &ep.bnf, nullptr, st));
    } else {
        MSX_TRY(conv_gather(d, true, x, w_packed, bias, y_raw, st));
        MSX_TRY(bn_eval_stats_launch(d->cout, running_mean, running_var, eps, save_mean, save_invstd, st));
    }
    return bn_apply_launch(d->dtype, y_raw, d->ldy, z, ldz, pixels, d->cout, save_mean, save_invstd, gamma, beta, relu, st);
}

int msx_conv_bn_act_bwd(const msx_conv_desc* d, const void* x, const float* w_master, const void* y_raw,
                        const void* dz, int32_t lddz, const float* gamma, const float* beta, const float* save_mean,
                        const float* save_invstd, int relu, void* w_packed, void* dy_scratch, void* dx, float* dw,
                        float* dbias, float* dgamma, float* dbeta, void* ws, size_t ws_bytes, void* stream) {
    MSX_TRY(validate_desc(d));
    MSX_REQUIRE(x && y_raw && dz && gamma && beta && save_mean && save_invstd && w_packed && dy_scratch,
                "conv_bn_act_bwd: null pointer");
    if (ws_bytes < ws_conv_bytes(d)) return set_error(MSX_ERR_WORKSPACE, "conv_bn_act_bwd: workspace %zu < %zu", ws_bytes, ws_conv_bytes(d));
    cudaStream_t st = as_stream(stream);
    const int64_t pixels = (int64_t)d->n * d->oh * d->ow;
    MSX_TRY(bn_bwd_launch(d->dtype, y_raw, d->ldy, dz, lddz, pixels, d->cout, gamma, beta, save_mean, save_invstd, relu,
                          dy_scratch, d->ldy, dgamma, dbeta, ws, ws_bytes, st));
    if (dx) {
        if (w_master) MSX_TRY(conv_pack_launch(d, MSX_PACK_DGRAD, w_master, w_packed, st));
        MSX_TRY(conv_gather(d, false, dy_scratch, w_packed, nullptr, dx, st));
    }
    if (dw) MSX_TRY(conv_wgrad(d, x, dy_scratch, dw, ws, ws_bytes, st));
    if (dbias) MSX_TRY(channel_sum_launch(d->dtype, dy_scratch, d->ldy, pixels, d->cout, dbias, ws, ws_bytes, st));
    return MSX_OK;
}

}  // extern "C"
