// seg_head: bias-free 1x1 convolution base_channel -> num_class at full resolution
// (models/unet.py:27, models/ducknet.py:35).  N_gemm = num_class is 1-2, so this is a pure
// bandwidth op: NHWC activations in, NCHW fp32 logits out (the layout the loss/metric kernels and
// the reference API use).  No tensor cores.
#include "common.cuh"

namespace msx {
namespace {
constexpr int kMaxClass = 8;

template <typename T>
__global__ void __launch_bounds__(256) head_fwd_kernel(const T* __restrict__ x, int ldx, int64_t hw, int64_t pixels, int cp, int channels,
                                                       const float* __restrict__ w, int nc, float* __restrict__ logits) {
    extern __shared__ float ws[];   // [nc][cp], zero padded
    for (int i = threadIdx.x; i < nc * cp; i += blockDim.x) {
        const int k = i / cp, c = i % cp;
        ws[i] = c < channels ? w[k * channels + c] : 0.f;
    }
    __syncthreads();
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < pixels; p += (int64_t)gridDim.x * blockDim.x) {
        float acc[kMaxClass];
#pragma unroll
        for (int k = 0; k < kMaxClass; ++k) acc[k] = 0.f;
        const T* xp = x + p * ldx;
        for (int c0 = 0; c0 < cp; c0 += 4) {
            float v[4];
            load4(xp + c0, v);
#pragma unroll
            for (int k = 0; k < kMaxClass; ++k)
                if (k < nc) {
                    const float* wk = ws + k * cp + c0;
                    acc[k] = fmaf(v[0], wk[0], fmaf(v[1], wk[1], fmaf(v[2], wk[2], fmaf(v[3], wk[3], acc[k]))));
                }
        }
        const int64_t n = p / hw, q = p % hw;
#pragma unroll
        for (int k = 0; k < kMaxClass; ++k)
            if (k < nc) logits[(n * nc + k) * hw + q] = acc[k];
    }
}

// Evaluation without logits in HBM (core/seg_trainer.py:114-119: preds = ema(images); metric.update(preds, masks)):
// the 1x1 head, torch.argmax (first maximal index, NaN maximal) and the confusion-matrix count in one pass over the
// features.  The logits are formed by exactly the fmaf chain of head_fwd_kernel, so every arg-max - ties included -
// and therefore every count equals what msx_head_fwd + msx_confmat_update produce.
template <typename T>
__global__ void __launch_bounds__(256) head_confmat_kernel(const T* __restrict__ x, int ldx, int64_t pixels, int cp, int channels,
                                                           const float* __restrict__ w, int nc, const int64_t* __restrict__ labels,
                                                           int64_t ignore, unsigned long long* __restrict__ confmat, int* __restrict__ bad) {
    extern __shared__ float ws[];                       // [nc][cp] weights, then counters
    unsigned int* cnt = reinterpret_cast<unsigned int*>(ws + nc * cp);   // nc <= 4: [nc*nc][256] private; else [nc*nc] shared atomics
    const bool priv = nc <= 4;
    const int cells = nc * nc;
    for (int i = threadIdx.x; i < nc * cp; i += blockDim.x) {
        const int k = i / cp, c = i % cp;
        ws[i] = c < channels ? w[k * channels + c] : 0.f;
    }
    if (priv) { for (int k = 0; k < cells; ++k) cnt[k * 256 + threadIdx.x] = 0; }
    else { for (int k = threadIdx.x; k < cells; k += blockDim.x) cnt[k] = 0; }
    __syncthreads();
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < pixels; p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = labels[p];
        float acc[kMaxClass];
#pragma unroll
        for (int k = 0; k < kMaxClass; ++k) acc[k] = 0.f;
        const T* xp = x + p * ldx;
        for (int c0 = 0; c0 < cp; c0 += 4) {
            float v[4];
            load4(xp + c0, v);
#pragma unroll
            for (int k = 0; k < kMaxClass; ++k)
                if (k < nc) {
                    const float* wk = ws + k * cp + c0;
                    acc[k] = fmaf(v[0], wk[0], fmaf(v[1], wk[1], fmaf(v[2], wk[2], fmaf(v[3], wk[3], acc[k]))));
                }
        }
        float best = acc[0];
        int bi = 0;
#pragma unroll
        for (int k = 1; k < kMaxClass; ++k)
            if (k < nc) {
                const bool take = (acc[k] > best) || (acc[k] != acc[k] && best == best);
                if (take) { best = acc[k]; bi = k; }
            }
        if (t == ignore) continue;
        if (t < 0 || t >= nc) { *bad = 1; continue; }
        const int cell = (int)t * nc + bi;
        if (priv) cnt[cell * 256 + threadIdx.x] += 1; else atomicAdd(&cnt[cell], 1u);
    }
    __syncthreads();
    if (priv) {
        for (int k = 0; k < cells; ++k) {
            unsigned int v = cnt[k * 256 + threadIdx.x];
            v = warp_sum(v);
            if ((threadIdx.x & 31) == 0 && v) atomicAdd(&confmat[k], (unsigned long long)v);
        }
    } else {
        for (int k = threadIdx.x; k < cells; k += blockDim.x)
            if (cnt[k]) atomicAdd(&confmat[k], (unsigned long long)cnt[k]);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) head_bwd_data_kernel(const float* __restrict__ dl, int64_t hw, int64_t pixels, int cp, int channels,
                                                            const float* __restrict__ w, int nc, T* __restrict__ dx, int lddx) {
    extern __shared__ float ws[];
    for (int i = threadIdx.x; i < nc * cp; i += blockDim.x) {
        const int k = i / cp, c = i % cp;
        ws[i] = c < channels ? w[k * channels + c] : 0.f;
    }
    __syncthreads();
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < pixels; p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t n = p / hw, q = p % hw;
        float g[kMaxClass];
#pragma unroll
        for (int k = 0; k < kMaxClass; ++k) g[k] = k < nc ? dl[(n * nc + k) * hw + q] : 0.f;
        T* op = dx + p * lddx;
        for (int c0 = 0; c0 < cp; c0 += 4) {
            float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int k = 0; k < kMaxClass; ++k)
                if (k < nc) {
                    const float* wk = ws + k * cp + c0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) v[j] = fmaf(g[k], wk[j], v[j]);
                }
            store4(op + c0, v);
        }
    }
}

// dw[k][c] = sum_p dl[k][p] * x[p][c]; channel-fixed threads (4 channels each), double partials
template <typename T>
__global__ void __launch_bounds__(256) head_bwd_weight_kernel(const T* __restrict__ x, int ldx, const float* __restrict__ dl, int64_t hw,
                                                              int64_t pixels, int cp, int vch, int nc, double* __restrict__ partials) {
    __shared__ double sh[256];
    const int lanes = blockDim.x / vch, lane = threadIdx.x / vch;
    const int c0 = (blockIdx.y * vch + threadIdx.x % vch) * 4;
    const bool active = lane < lanes && c0 < cp;
    float acc[kMaxClass][4];
#pragma unroll
    for (int k = 0; k < kMaxClass; ++k)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[k][j] = 0.f;
    if (active) {
        // (image, offset) of the pixel advanced incrementally: the 64-bit p / hw, p % hw per pixel made this
        // kernel 20x slower than its traffic
        const int64_t step = (int64_t)gridDim.x * lanes;
        int64_t p = (int64_t)blockIdx.x * lanes + lane;
        int64_t n = p / hw, q = p - n * hw;
        auto advance = [&](int64_t& nn, int64_t& qq) { qq += step; while (qq >= hw) { qq -= hw; ++nn; } };
        for (; p + step < pixels; p += 2 * step) {                      // two independent pixels in flight
            int64_t n1 = n, q1 = q;
            advance(n1, q1);
            float v0[4], v1[4], g0[kMaxClass], g1[kMaxClass];
            load4(x + p * ldx + c0, v0);
            load4(x + (p + step) * ldx + c0, v1);
#pragma unroll
            for (int k = 0; k < kMaxClass; ++k)
                if (k < nc) { g0[k] = dl[(n * nc + k) * hw + q]; g1[k] = dl[(n1 * nc + k) * hw + q1]; }
#pragma unroll
            for (int k = 0; k < kMaxClass; ++k)
                if (k < nc) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[k][j] = fmaf(g1[k], v1[j], fmaf(g0[k], v0[j], acc[k][j]));
                }
            n = n1; q = q1;
            advance(n, q);
        }
        if (p < pixels) {
            float v[4];
            load4(x + p * ldx + c0, v);
#pragma unroll
            for (int k = 0; k < kMaxClass; ++k)
                if (k < nc) {
                    const float g = dl[(n * nc + k) * hw + q];
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[k][j] = fmaf(g, v[j], acc[k][j]);
                }
        }
    }
    int p2 = 1;
    while (p2 < lanes) p2 <<= 1;
#pragma unroll
    for (int k = 0; k < kMaxClass; ++k) {
        if (k >= nc) break;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            sh[threadIdx.x] = (double)acc[k][j];
            __syncthreads();
            for (int off = p2 >> 1; off > 0; off >>= 1) {
                if (lane < off && lane + off < lanes) sh[threadIdx.x] += sh[threadIdx.x + off * vch];
                __syncthreads();
            }
            if (active && lane == 0) partials[((int64_t)blockIdx.x * nc + k) * cp + c0 + j] = sh[threadIdx.x];
            __syncthreads();
        }
    }
}
__global__ void head_bwd_weight_finalize_kernel(const double* __restrict__ partials, int nblk, int nc, int cp, int channels, float* __restrict__ dw) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nc * channels) return;
    const int k = i / channels, c = i % channels;
    double s = 0.0;
    for (int b = 0; b < nblk; ++b) s += partials[((int64_t)b * nc + k) * cp + c];
    dw[i] = (float)s;
}
}  // namespace
}  // namespace msx

using namespace msx;

extern "C" {

int msx_head_fwd(int dtype, const void* x, int32_t ldx, int32_t n, int32_t h, int32_t w, int32_t channels,
                 const float* w_master, int32_t num_class, float* logits, void* stream) {
    MSX_REQUIRE(x && w_master && logits, "head_fwd: null pointer");
    MSX_REQUIRE(num_class >= 1 && num_class <= kMaxClass, "head: num_class %d not in [1,%d]", num_class, kMaxClass);
    const int cp = round_up(channels, 8);
    MSX_REQUIRE(ldx % 8 == 0 && ldx >= cp, "head_fwd: bad ldx %d", ldx);
    const size_t smem = (size_t)num_class * cp * sizeof(float);
    MSX_REQUIRE(smem <= 48 * 1024, "head: num_class*channels too large for shared memory");
    const int64_t hw = (int64_t)h * w, pixels = hw * n;
    const int grid = (int)(cdiv64(pixels, 256) < kNumSMs * 8 ? cdiv64(pixels, 256) : kNumSMs * 8);
    MSX_DISPATCH_DTYPE(dtype, T, {
        { ProfScope ps__(K_HEAD, as_stream(stream), (double)pixels * (channels * (double)sizeof(T) + num_class * 4.0), 2.0 * (double)pixels * channels * num_class); head_fwd_kernel<T><<<grid, 256, smem, as_stream(stream)>>>(static_cast<const T*>(x), ldx, hw, pixels, cp, channels, w_master, num_class, logits); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_head_confmat_update(int dtype, const void* x, int32_t ldx, int64_t pixels, int32_t channels, const float* w_master,
                            int32_t num_class, const int64_t* labels, int64_t ignore_index, int64_t* confmat,
                            int32_t* bad_label_flag, void* stream) {
    MSX_REQUIRE(x && w_master && labels && confmat && bad_label_flag, "head_confmat: null pointer");
    MSX_REQUIRE(num_class >= 1 && num_class <= kMaxClass, "head: num_class %d not in [1,%d]", num_class, kMaxClass);
    const int cp = round_up(channels, 8);
    MSX_REQUIRE(ldx % 8 == 0 && ldx >= cp, "head_confmat: bad ldx %d", ldx);
    const int cells = num_class * num_class;
    const size_t smem = (size_t)num_class * cp * sizeof(float) + (num_class <= 4 ? (size_t)cells * 256 : (size_t)cells) * sizeof(unsigned int);
    MSX_REQUIRE(smem <= 48 * 1024, "head_confmat: num_class*channels too large for shared memory");
    const int grid = (int)(cdiv64(pixels, 256) < kNumSMs * 8 ? cdiv64(pixels, 256) : kNumSMs * 8);
    auto* cm = reinterpret_cast<unsigned long long*>(confmat);
    MSX_DISPATCH_DTYPE(dtype, T, {
        { ProfScope ps__(K_CONFMAT, as_stream(stream), (double)pixels * (channels * (double)sizeof(T) + 8.0), 2.0 * (double)pixels * channels * num_class);
          head_confmat_kernel<T><<<grid, 256, smem, as_stream(stream)>>>(static_cast<const T*>(x), ldx, pixels, cp, channels, w_master, num_class, labels, ignore_index, cm, bad_label_flag); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_head_bwd(int dtype, const void* x, int32_t ldx, int32_t n, int32_t h, int32_t w, int32_t channels,
                 const float* w_master, int32_t num_class, const float* dlogits, void* dx, int32_t lddx, float* dw,
                 void* ws, size_t ws_bytes, void* stream) {
    MSX_REQUIRE(x && w_master && dlogits, "head_bwd: null pointer");
    MSX_REQUIRE(num_class >= 1 && num_class <= kMaxClass, "head: num_class %d not in [1,%d]", num_class, kMaxClass);
    const int cp = round_up(channels, 8);
    const size_t smem = (size_t)num_class * cp * sizeof(float);
    MSX_REQUIRE(smem <= 48 * 1024, "head: num_class*channels too large for shared memory");
    const int64_t hw = (int64_t)h * w, pixels = hw * n;
    cudaStream_t st = as_stream(stream);
    if (dx) {
        MSX_REQUIRE(lddx % 8 == 0 && lddx >= cp, "head_bwd: bad lddx %d", lddx);
        const int grid = (int)(cdiv64(pixels, 256) < kNumSMs * 8 ? cdiv64(pixels, 256) : kNumSMs * 8);
        MSX_DISPATCH_DTYPE(dtype, T, {
            { ProfScope ps__(K_HEAD, st, (double)pixels * (channels * (double)sizeof(T) + num_class * 4.0), 2.0 * (double)pixels * channels * num_class); head_bwd_data_kernel<T><<<grid, 256, smem, st>>>(dlogits, hw, pixels, cp, channels, w_master, num_class, static_cast<T*>(dx), lddx); }
        });
        MSX_LAUNCH_CHECK();
    }
    if (dw) {
        if (ws_bytes < bn_partials_bytes(channels)) return set_error(MSX_ERR_WORKSPACE, "head_bwd: workspace too small");
        const int V = cp / 4, vch = V < 64 ? V : 64, nchunk = (V + vch - 1) / vch, lanes = 256 / vch;
        int64_t g = cdiv64(pixels, (int64_t)lanes * 8);
        const int64_t gmax = 2 * kMaxPartialBlocks / num_class;     // partials [g][num_class][cp] must fit bn_partials_bytes
        if (g > kNumSMs * 4) g = kNumSMs * 4;
        if (g > gmax) g = gmax;
        if (g < 1) g = 1;
        MSX_DISPATCH_DTYPE(dtype, T, {
            { ProfScope ps__(K_HEAD, st, (double)pixels * (channels * (double)sizeof(T) + num_class * 4.0), 2.0 * (double)pixels * channels * num_class); head_bwd_weight_kernel<T><<<dim3((unsigned)g, nchunk), 256, 0, st>>>(static_cast<const T*>(x), ldx, dlogits, hw, pixels, cp, vch,
                                                                                 num_class, static_cast<double*>(ws)); }
        });
        { ProfScope ps__(K_MISC, st, 0, 0); head_bwd_weight_finalize_kernel<<<(num_class * channels + 127) / 128, 128, 0, st>>>(static_cast<double*>(ws), (int)g, num_class, cp, channels, dw); }
        MSX_LAUNCH_CHECK();
    }
    return MSX_OK;
}

}  // extern "C"
