// Shared helpers for libmsx.so (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/msx.h"

namespace msx {

int set_error(int code, const char* fmt, ...);
void note_conv_engine(int engine);

#define MSX_CHECK_CUDA(expr)                                                                       \
    do {                                                                                           \
        cudaError_t e__ = (expr);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return msx::set_error(MSX_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr,       \
                                  cudaGetErrorString(e__));                                        \
    } while (0)
#define MSX_REQUIRE(cond, ...)                                                                     \
    do {                                                                                           \
        if (!(cond)) return msx::set_error(MSX_ERR_INVALID, __VA_ARGS__);                          \
    } while (0)
#define MSX_LAUNCH_CHECK() MSX_CHECK_CUDA(cudaGetLastError())
#define MSX_TRY(expr)                                                                              \
    do {                                                                                           \
        int rc__ = (expr);                                                                         \
        if (rc__ != MSX_OK) return rc__;                                                           \
    } while (0)

constexpr int kNumSMs = 148;            // B200
constexpr int kMaxPartialBlocks = 592;  // 4 CTAs per SM for the channel-reduction kernels

__host__ __device__ inline int round_up(int x, int m) { return (x + m - 1) / m * m; }
__host__ __device__ inline int64_t cdiv64(int64_t a, int64_t b) { return (a + b - 1) / b; }

typedef __nv_bfloat16 bf16;

// 16-byte vectors of activations: 4 floats or 8 bf16.
template <typename T>
struct Vec;
template <>
struct Vec<float> {
    static constexpr int N = 4;
    __device__ static inline void load(const float* p, float (&v)[4]) {
        float4 r = *reinterpret_cast<const float4*>(p);
        v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w;
    }
    __device__ static inline void store(float* p, const float (&v)[4]) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <>
struct Vec<bf16> {
    static constexpr int N = 8;
    __device__ static inline void load(const bf16* p, float (&v)[8]) {
        uint4 r = *reinterpret_cast<const uint4*>(p);
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            v[2 * i] = __uint_as_float(w[i] << 16);
            v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    __device__ static inline void store(bf16* p, const float (&v)[8]) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
};

// 4-element access used by the SIMT conv engine (16 B of float / 8 B of bf16).
__device__ inline void load4(const float* p, float (&v)[4]) { Vec<float>::load(p, v); }
__device__ inline void load4(const bf16* p, float (&v)[4]) {
    uint2 r = *reinterpret_cast<const uint2*>(p);
    v[0] = __uint_as_float(r.x << 16); v[1] = __uint_as_float(r.x & 0xffff0000u);
    v[2] = __uint_as_float(r.y << 16); v[3] = __uint_as_float(r.y & 0xffff0000u);
}
__device__ inline void store4(float* p, const float (&v)[4]) { Vec<float>::store(p, v); }
__device__ inline void store4(bf16* p, const float (&v)[4]) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
    __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
    *reinterpret_cast<uint2*>(p) = make_uint2(*reinterpret_cast<uint32_t*>(&a), *reinterpret_cast<uint32_t*>(&b));
}
__device__ inline void store2(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
__device__ inline void store2(bf16* p, float a, float b) {
    *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b);
}
__device__ inline float to_float(float x) { return x; }
__device__ inline float to_float(bf16 x) { return __bfloat162float(x); }
template <typename T> __device__ inline T from_float(float x);
template <> __device__ inline float from_float<float>(float x) { return x; }
template <> __device__ inline bf16 from_float<bf16>(float x) { return __float2bfloat16_rn(x); }

// Warp / block sums (double for the statistics paths).
template <typename T>
__device__ inline T warp_sum(T v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// Block-wide sum; result valid in thread 0.  `sh` needs >= 32 entries.
template <typename T>
__device__ inline T block_sum(T v, T* sh) {
    v = warp_sum(v);
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    __syncthreads();
    if (lane == 0) sh[wid] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    v = (threadIdx.x < nw) ? sh[threadIdx.x] : T(0);
    if (wid == 0) v = warp_sum(v);
    return v;
}

// ---- launch accounting / live profiler (api.cu) -----------------------------------------------
// Every kernel launch sits inside a ProfScope: it always bumps the library-wide launch counter
// (bench.py's gpu_launches) and, when profiling is enabled (msx_profile_enable), brackets the launch
// with CUDA events on the launching stream and books the launch's ALGORITHMIC bytes / flops, so
// that bench.py can form achieved GB/s / TFLOP/s per kernel class from live timings.
enum KClass {
    K_CONV_SIMT = 0, K_WGRAD_SIMT, K_WGRAD_REDUCE, K_PACK, K_BN_STATS, K_BN_FINALIZE, K_BN_APPLY, K_BN_BWD_REDUCE,
    K_BN_BWD_APPLY, K_SUM, K_UPSAMPLE_BWD, K_MAXPOOL, K_LAYOUT, K_HEAD, K_LOSS, K_CONFMAT, K_OPTIM, K_CONV_TC,
    K_WGRAD_TC, K_MISC, K_NUM
};
struct ProfScope {
    int cls; cudaStream_t st; void* rec;
    ProfScope(int cls, cudaStream_t st, double bytes, double flops);
    ~ProfScope();
};

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// dtype dispatch
#define MSX_DISPATCH_DTYPE(dtype, T, ...)                                          \
    do {                                                                           \
        if ((dtype) == MSX_F32) { using T = float; __VA_ARGS__; }                  \
        else if ((dtype) == MSX_BF16) { using T = msx::bf16; __VA_ARGS__; }        \
        else return msx::set_error(MSX_ERR_INVALID, "unknown dtype %d", (int)(dtype)); \
    } while (0)

// What a gather-form conv kernel does to an accumulator column c before storing it:
//   v += bias[c] (ConvTranspose2d bias);  v = v * scale[c] + shift[c] (eval-mode BatchNorm folded to an affine);  ReLU.
// All pointers nullable; columns >= the logical channel count store zeros.
struct ConvEpilogue {
    const float* bias;
    const float* scale;
    const float* shift;
    int relu;
};

// ---- internal cross-file entry points -------------------------------------------------------
// channel-reduction partial sums: [nblk][2][cp] doubles
size_t bn_partials_bytes(int channels);
int bn_stats_launch(int dtype, const void* x, int ld, int64_t pixels, int channels, double* partials,
                    int* nblk_out, cudaStream_t st);
int bn_finalize_launch(const double* partials, int nblk, int channels, int64_t count, float eps, float momentum,
                       float* mean, float* invstd, float* running_mean, float* running_var,
                       int64_t* num_batches_tracked, cudaStream_t st);
int bn_eval_stats_launch(int channels, const float* running_mean, const float* running_var, float eps,
                         float* mean, float* invstd, cudaStream_t st);
// eval-mode BN as y*scale + shift (scale = gamma*invstd, shift = beta - mean*scale [+ bias*scale]); arrays of round_up(channels, 8)
int bn_eval_affine_launch(int channels, const float* running_mean, const float* running_var, float eps,
                          const float* gamma, const float* beta, const float* conv_bias,
                          float* mean, float* invstd, float* scale, float* shift, cudaStream_t st);
int bn_apply_launch(int dtype, const void* x, int ldx, void* z, int ldz, int64_t pixels, int channels,
                    const float* mean, const float* invstd, const float* gamma, const float* beta, int relu,
                    cudaStream_t st);
int bn_bwd_launch(int dtype, const void* x, int ldx, const void* dz, int lddz, int64_t pixels, int channels,
                  const float* gamma, const float* beta, const float* mean, const float* invstd, int flags /* MSX_BN_* */,
                  void* dx, int lddx, float* dgamma, float* dbeta, void* ws, size_t ws_bytes, cudaStream_t st);
int channel_sum_launch(int dtype, const void* x, int ld, int64_t pixels, int channels, float* out,
                       void* ws, size_t ws_bytes, cudaStream_t st);

}  // namespace msx
