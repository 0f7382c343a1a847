// Flat-arena optimizer / EMA kernels: one launch per step instead of one per tensor
// (utils/optimizer.py -> torch.optim.{Adam,AdamW,SGD}; utils/model_ema.py:29-40 copies 1733 tensors
// per DUCKNet step).  Pure streaming: Adam 28 B/param, SGD-momentum 20 B/param, EMA 12 B/param.
#include "common.cuh"

namespace msx {
namespace {
__global__ void step_inc_kernel(int64_t* step) { *step += 1; }

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, int64_t count, float lr, float b1, float b2, float eps,
                                                   float wd, int decoupled, float gscale, const int64_t* __restrict__ step,
                                                   const float* __restrict__ hyper) {
    if (hyper) { lr = hyper[0]; b1 = hyper[1]; b2 = hyper[2]; wd = hyper[3]; }
    const double t = (double)*step;
    const float bc1 = (float)(1.0 - pow((double)b1, t));
    const float bc2s = (float)sqrt(1.0 - pow((double)b2, t));
    const float step_size = lr / bc1;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        float pi = p[i], gi = g[i] * gscale;
        if (decoupled) pi *= 1.f - lr * wd; else if (wd != 0.f) gi += wd * pi;
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        m[i] = mi; v[i] = vi;
        p[i] = pi - step_size * (mi / (sqrtf(vi) / bc2s + eps));
    }
}
__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ buf,
                                                  int64_t count, float lr, float mom, float wd, float gscale, int first,
                                                  const float* __restrict__ hyper) {
    if (hyper) { lr = hyper[0]; mom = hyper[1]; wd = hyper[3]; }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const float pi = p[i];
        float gi = g[i] * gscale + wd * pi;
        if (mom != 0.f) {
            const float b = first ? gi : mom * buf[i] + gi;
            buf[i] = b;
            gi = b;
        }
        p[i] = pi - lr * gi;
    }
}
// decay * e + (1 - decay) * m exactly as the reference's three torch kernels round it: no FMA contraction
__global__ void __launch_bounds__(256) ema_kernel(float* __restrict__ e, const float* __restrict__ m, int64_t count, float decay, float omd, int copy) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
        e[i] = copy ? m[i] : __fadd_rn(__fmul_rn(decay, e[i]), __fmul_rn(omd, m[i]));
}
__global__ void __launch_bounds__(256) ema_i64_kernel(int64_t* __restrict__ e, const int64_t* __restrict__ m, int64_t count, float decay, float omd, int copy) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
        e[i] = copy ? m[i] : (int64_t)__fadd_rn(__fmul_rn(decay, (float)e[i]), __fmul_rn(omd, (float)m[i]));   // copy_ float -> int64 truncates
}
inline int flat_grid(int64_t count) {
    int64_t g = cdiv64(count, 256 * 4);
    if (g < 1) g = 1;
    if (g > kNumSMs * 8) g = kNumSMs * 8;
    return (int)g;
}
}  // namespace
}  // namespace msx

using namespace msx;

extern "C" {
int msx_adam_step(float* param, const float* grad, float* exp_avg, float* exp_avg_sq, int64_t count, float lr, float beta1,
                  float beta2, float eps, float weight_decay, int decoupled_wd, float grad_scale, int64_t* step_dev, const float* hyper_dev, void* stream) {
    MSX_REQUIRE(param && grad && exp_avg && exp_avg_sq && step_dev, "adam: null pointer");
    cudaStream_t st = as_stream(stream);
    { ProfScope ps__(K_MISC, st, 0, 0); step_inc_kernel<<<1, 1, 0, st>>>(step_dev); }
    { ProfScope ps__(K_OPTIM, st, (double)count * 28.0, 0); adam_kernel<<<flat_grid(count), 256, 0, st>>>(param, grad, exp_avg, exp_avg_sq, count, lr, beta1, beta2, eps, weight_decay,
                                                  decoupled_wd, grad_scale, step_dev, hyper_dev); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}
int msx_sgd_step(float* param, const float* grad, float* momentum_buf, int64_t count, float lr, float momentum,
                 float weight_decay, float grad_scale, int first_step, const float* hyper_dev, void* stream) {
    MSX_REQUIRE(param && grad && (momentum == 0.f || momentum_buf), "sgd: null pointer");
    { ProfScope ps__(K_OPTIM, as_stream(stream), (double)count * 20.0, 0); sgd_kernel<<<flat_grid(count), 256, 0, as_stream(stream)>>>(param, grad, momentum_buf, count, lr, momentum, weight_decay, grad_scale, first_step, hyper_dev); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}
int msx_ema_update(float* ema, const float* model, int64_t count, double decay, void* stream) {
    MSX_REQUIRE(ema && model, "ema: null pointer");
    // the reference multiplies by the Python doubles `decay` and `1. - decay`, each rounded to fp32 by the torch kernel
    { ProfScope ps__(K_OPTIM, as_stream(stream), (double)count * (decay == 0.0 ? 8.0 : 12.0), 0); ema_kernel<<<flat_grid(count), 256, 0, as_stream(stream)>>>(ema, model, count, (float)decay, (float)(1.0 - decay), decay == 0.0); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}
int msx_ema_update_i64(int64_t* ema, const int64_t* model, int64_t count, double decay, void* stream) {
    MSX_REQUIRE(ema && model, "ema_i64: null pointer");
    { ProfScope ps__(K_OPTIM, as_stream(stream), (double)count * 24.0, 0); ema_i64_kernel<<<flat_grid(count), 256, 0, as_stream(stream)>>>(ema, model, count, (float)decay, (float)(1.0 - decay), decay == 0.0); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}
}
