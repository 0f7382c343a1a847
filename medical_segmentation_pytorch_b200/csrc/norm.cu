// Bandwidth-bound NHWC kernels: BatchNorm statistics / apply / backward, n-way sum (+nearest-2x
// upsample) with fused statistics, MaxPool2d(3,2,1), layout conversion.
//
// Thread mapping shared by all of them ("channel-fixed"): a thread owns one 16-byte channel
// vector (4 floats / 8 bf16) and walks over pixels, so per-channel parameters and accumulators
// stay in registers and every warp access is a run of whole 16 B vectors along the contiguous
// channel axis.  A CTA of 256 threads = `lanes` pixels x `vch` vectors; blockIdx.y selects a
// chunk of <= 64 vectors when the tensor has more channels than that.
#include <stdlib.h>

#include <map>
#include <mutex>
#include <utility>

#include "common.cuh"

namespace msx {

namespace {

struct ChanCfg {
    int cp, vch, nchunk, lanes;
};
template <int VEC>
ChanCfg chan_cfg(int channels) {
    ChanCfg c;
    c.cp = round_up(channels, 8);
    const int V = c.cp / VEC;
    c.vch = V < 64 ? V : 64;
    c.nchunk = (V + c.vch - 1) / c.vch;
    c.lanes = 256 / c.vch;
    return c;
}

struct ChanThread {
    int c0, lane, lanes;
    bool active;
};
template <int VEC>
__device__ inline ChanThread chan_thread(int cp, int vch) {
    ChanThread t;
    t.lanes = blockDim.x / vch;
    t.lane = threadIdx.x / vch;
    const int v = blockIdx.y * vch + threadIdx.x % vch;
    t.c0 = v * VEC;
    t.active = t.lane < t.lanes && t.c0 < cp;
    return t;
}

// Sum v[j] over the pixel lanes of the CTA (threads that share threadIdx.x % vch); result valid
// in lane 0.  sh: NV * 256 doubles; all NV values go through ONE tree (log2(lanes) barriers in total).
#ifndef MSX_BN_MINBLOCKS
#define MSX_BN_MINBLOCKS 1          // CTAs per SM the reduction kernels are compiled for (register cap)
#endif
#ifdef MSX_BN_SMALL_SMEM
constexpr int kLaneReduceRounds = 2;   // half the shared memory (16 KB for bf16): the values go through the tree in two rounds
#else
constexpr int kLaneReduceRounds = 1;
#endif
template <int NV>
__device__ inline void lane_reduce(double (&v)[NV], double* sh, int vch, int lane, int lanes) {
    int p2 = 1;
    while (p2 < lanes) p2 <<= 1;
    constexpr int H = NV / kLaneReduceRounds;
#pragma unroll
    for (int r = 0; r < kLaneReduceRounds; ++r) {
        if (r) __syncthreads();
#pragma unroll
        for (int j = 0; j < H; ++j) sh[j * 256 + threadIdx.x] = v[r * H + j];
        __syncthreads();
        for (int off = p2 >> 1; off > 0; off >>= 1) {
            if (lane < off && lane + off < lanes) {
#pragma unroll
                for (int j = 0; j < H; ++j) sh[j * 256 + threadIdx.x] += sh[j * 256 + threadIdx.x + off * vch];
            }
            __syncthreads();
        }
#pragma unroll
        for (int j = 0; j < H; ++j) v[r * H + j] = sh[j * 256 + threadIdx.x];
    }
}

inline int reduce_grid(int64_t pixels, int lanes) {
    int64_t g = cdiv64(pixels, (int64_t)lanes * 8);
    if (g < 1) g = 1;
    if (g > kNumSMs * 2) g = kNumSMs * 2;      // one resident wave (2 CTAs/SM by registers); halves the finalize work
    return (int)g;
}
inline int stream_grid(int64_t pixels, int lanes) {
    int64_t g = cdiv64(pixels, (int64_t)lanes * 4);
    if (g < 1) g = 1;
    if (g > kNumSMs * 6) g = kNumSMs * 6;      // few, fat CTAs: the per-CTA parameter prologue is amortised
    return (int)g;
}

// ---- statistics ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256, MSX_BN_MINBLOCKS) bn_stats_kernel(const T* __restrict__ x, int ld, int64_t pixels, int cp, int vch,
                                                       double* partials, const BnFinalize fin) {
    constexpr int VEC = Vec<T>::N;
    __shared__ double sh[2 * VEC * 256 / kLaneReduceRounds];
    __shared__ int last_flag;
    const ChanThread t = chan_thread<VEC>(cp, vch);
    double acc[2 * VEC];
#pragma unroll
    for (int j = 0; j < 2 * VEC; ++j) acc[j] = 0.0;
    if (t.active) {
        const int64_t step = (int64_t)gridDim.x * t.lanes;
        int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane;
        for (; p + 3 * step < pixels; p += 4 * step) {           // 4 independent 16 B loads in flight per thread
            float v[4][VEC];
#pragma unroll
            for (int u = 0; u < 4; ++u) Vec<T>::load(x + (p + u * step) * ld + t.c0, v[u]);
            // fp32 over the 4 pixels, ONE F2F.F64 per channel and sum: the fp32->fp64 convert pipe, not HBM,
            // was the limiter when every element was converted
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                const float fs = (v[0][j] + v[1][j]) + (v[2][j] + v[3][j]);
                const float fq = fmaf(v[0][j], v[0][j], v[1][j] * v[1][j]) + fmaf(v[2][j], v[2][j], v[3][j] * v[3][j]);
                acc[j] += (double)fs;
                acc[VEC + j] += (double)fq;
            }
        }
        for (; p < pixels; p += step) {
            float v[VEC];
            Vec<T>::load(x + p * ld + t.c0, v);
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                acc[j] += (double)v[j];
                acc[VEC + j] += (double)v[j] * (double)v[j];
            }
        }
    }
    lane_reduce<2 * VEC>(acc, sh, vch, t.lane, t.lanes);
    if (t.active && t.lane == 0) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            partials[((int64_t)blockIdx.x * 2 + 0) * cp + t.c0 + j] = acc[j];
            partials[((int64_t)blockIdx.x * 2 + 1) * cp + t.c0 + j] = acc[VEC + j];
        }
    }
    if (fin.ticket && ticket_is_last(fin.ticket, gridDim.x * gridDim.y, &last_flag)) bn_finalize_all(fin, gridDim.x);
}

// Fold of one channel by a whole 128-thread block (fixed order): every thread takes rows tid, tid + 128, ... (at most
// four independent loads for <= 592 rows), then warp shuffles and one pass over the four warp sums.  One L2 round trip
// instead of the four to five a single warp needs - these kernels are pure latency on the step's critical path.
__device__ inline void fold_partials_block(const double* partials, int nblk, int cp, int c, double& s, double& q, double* sh) {
    double a = 0.0, b = 0.0;
    if (nblk <= 5 * 128) {                       // kMaxPartialBlocks = 592: all ten loads of a thread in flight at once
        double va[5], vb[5];
#pragma unroll
        for (int u = 0; u < 5; ++u) {
            const int r = threadIdx.x + u * 128;
            va[u] = r < nblk ? __ldcg(partials + ((int64_t)r * 2 + 0) * cp + c) : 0.0;
            vb[u] = r < nblk ? __ldcg(partials + ((int64_t)r * 2 + 1) * cp + c) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < 5; ++u) { a += va[u]; b += vb[u]; }
    } else {
        for (int r = threadIdx.x; r < nblk; r += 128) {
            a += __ldcg(partials + ((int64_t)r * 2 + 0) * cp + c);
            b += __ldcg(partials + ((int64_t)r * 2 + 1) * cp + c);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        a += __shfl_xor_sync(0xffffffffu, a, o);
        b += __shfl_xor_sync(0xffffffffu, b, o);
    }
    if ((threadIdx.x & 31) == 0) { sh[(threadIdx.x >> 5) * 2] = a; sh[(threadIdx.x >> 5) * 2 + 1] = b; }
    __syncthreads();
    s = (sh[0] + sh[2]) + (sh[4] + sh[6]);
    q = (sh[1] + sh[3]) + (sh[5] + sh[7]);
}

// Stand-alone fold + finalize: one 128-thread block per channel.  Measured (round 2): letting the reduction launch's last
// CTA do this instead (tickets, common.cuh) removes a launch but serialises the fold of all channels in ONE CTA, which
// costs more than the launch for every layer wider than ~40 channels (bn_bwd_reduce 6.8 -> 15.2 ms per DUCKNet-17 step),
// so the in-kernel path is only taken up to ticket_max_cp() channels (default 0: never).
__global__ void __launch_bounds__(128) bn_finalize_kernel(int nblk, const BnFinalize f) {
    __shared__ double sh[8];
    pdl_wait();
    pdl_trigger();
    const int c = blockIdx.x;
    const int lane = threadIdx.x;
    if (c == 0 && lane == 0 && f.nbt && !f.sums) *f.nbt += 1;
    if (c == 0 && lane == 0 && f.sums) f.sums[2 * f.cp] = (double)f.count;      // this rank's sample count, reduced with the sums
    if (c >= f.cp) return;
    if (c >= f.channels) {
        if (lane == 0) {
            if (f.sums) { f.sums[c] = 0.0; f.sums[f.cp + c] = 0.0; }
            else { f.mean[c] = 0.f; f.invstd[c] = 0.f; }
        }
        return;
    }
    double s, q;
    fold_partials_block(f.partials, nblk, f.cp, c, s, q, sh);
    if (lane != 0) return;
    if (f.sums) { f.sums[c] = s; f.sums[f.cp + c] = q; }
    else bn_finalize_one(f, c, s, q);
}

// SyncBatchNorm: per-channel [sum, sum of squares] (doubles, already reduced over ranks) -> mean / invstd / running stats
__global__ void bn_finalize_sums_kernel(const double* __restrict__ sums, BnFinalize f) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c == 0 && f.nbt) *f.nbt += 1;
    if (c >= f.cp) return;
    if (c >= f.channels) { f.mean[c] = 0.f; f.invstd[c] = 0.f; return; }
    f.count = (long long)(sums[2 * f.cp] + 0.5);               // samples over all ranks
    bn_finalize_one(f, c, sums[c], sums[f.cp + c]);
}

__global__ void bn_eval_stats_kernel(int cp, int channels, const float* rm, const float* rv, float eps,
                                     float* mean, float* invstd) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cp) return;
    if (c >= channels) { mean[c] = 0.f; invstd[c] = 0.f; return; }
    mean[c] = rm[c];
    invstd[c] = (float)(1.0 / sqrt((double)rv[c] + (double)eps));
}

__global__ void bn_eval_affine_kernel(int cp, int channels, const float* rm, const float* rv, float eps, const float* gamma,
                                      const float* beta, const float* conv_bias, float* mean, float* invstd, float* scale, float* shift) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cp) return;
    if (c >= channels) { mean[c] = 0.f; invstd[c] = 0.f; scale[c] = 0.f; shift[c] = 0.f; return; }
    const float is = (float)(1.0 / sqrt((double)rv[c] + (double)eps));
    const float sc = is * gamma[c];                       // the same product bn_apply forms
    mean[c] = rm[c];
    invstd[c] = is;
    scale[c] = sc;
    shift[c] = fmaf((conv_bias ? conv_bias[c] : 0.f) - rm[c], sc, beta[c]);
}

// ---- apply --------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) bn_apply_kernel(const T* __restrict__ x, int ldx, T* __restrict__ z, int ldz,
                                                       int64_t pixels, int cp, int vch, int channels,
                                                       const float* __restrict__ mean, const float* __restrict__ invstd,
                                                       const float* __restrict__ gamma, const float* __restrict__ beta,
                                                       int relu) {
    constexpr int VEC = Vec<T>::N;
    const ChanThread t = chan_thread<VEC>(cp, vch);
    if (!t.active) return;
    float mu[VEC], sc[VEC], sh[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        const int c = t.c0 + j;
        const bool ok = c < channels;
        mu[j] = ok ? mean[c] : 0.f;
        sc[j] = ok ? invstd[c] * gamma[c] : 0.f;
        sh[j] = ok ? beta[c] : 0.f;
    }
    const int64_t step = (int64_t)gridDim.x * t.lanes;
    int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane;
    for (; p + 3 * step < pixels; p += 4 * step) {
        float v[4][VEC];
#pragma unroll
        for (int u = 0; u < 4; ++u) Vec<T>::load(x + (p + u * step) * ldx + t.c0, v[u]);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                float r = (v[u][j] - mu[j]) * sc[j] + sh[j];
                v[u][j] = relu ? fmaxf(r, 0.f) : r;
            }
            Vec<T>::store(z + (p + u * step) * ldz + t.c0, v[u]);
        }
    }
    for (; p < pixels; p += step) {
        float v[VEC];
        Vec<T>::load(x + p * ldx + t.c0, v);
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            float r = (v[j] - mu[j]) * sc[j] + sh[j];
            v[j] = relu ? fmaxf(r, 0.f) : r;
        }
        Vec<T>::store(z + p * ldz + t.c0, v);
    }
}

// ---- backward -----------------------------------------------------------------------------
// What the backward sums become: dbeta = sum g, dgamma = sum g*xhat, and the two per-channel coefficients of
// dx = gamma*invstd*(g - c1 - xhat*c2): c1 = sum g / count, c2 = sum g*xhat / count (both 0 in eval mode, where mean and
// invstd are constants).  With `sums` (SyncBatchNorm) the two sums also go to [2][cp] doubles for the cross-rank
// reduction that msx_bn_bwd_coeffs_from_sums turns into c1 / c2.
struct BnBwdFinalize {
    unsigned int* ticket;
    double* sums;
    float *dgamma, *dbeta, *c1, *c2;
    long long count;
    int channels, cp, eval_mode;
};

__device__ inline void bn_bwd_finalize_all(const BnBwdFinalize& f, const double* partials, int nblk) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    for (int c = warp; c < f.cp; c += nwarps) {
        if (c >= f.channels) {
            if (lane == 0) {
                if (f.sums) { f.sums[c] = 0.0; f.sums[f.cp + c] = 0.0; }
                f.c1[c] = 0.f; f.c2[c] = 0.f;
            }
            continue;
        }
        double s, q;
        fold_partials(partials, nblk, f.cp, c, s, q);
        if (lane != 0) continue;
        if (f.dbeta) f.dbeta[c] = (float)s;
        if (f.dgamma) f.dgamma[c] = (float)q;
        if (f.sums) { f.sums[c] = s; f.sums[f.cp + c] = q; }
        f.c1[c] = f.eval_mode ? 0.f : (float)(s / (double)f.count);
        f.c2[c] = f.eval_mode ? 0.f : (float)(q / (double)f.count);
    }
}

__global__ void __launch_bounds__(128) bn_bwd_finalize_kernel(const double* partials, int nblk, const BnBwdFinalize f) {
    __shared__ double sh[8];
    pdl_wait();
    pdl_trigger();
    const int c = blockIdx.x;
    const int lane = threadIdx.x;
    if (c >= f.cp) return;
    if (c >= f.channels) {
        if (lane == 0) {
            if (f.sums) { f.sums[c] = 0.0; f.sums[f.cp + c] = 0.0; }
            f.c1[c] = 0.f; f.c2[c] = 0.f;
        }
        return;
    }
    double s, q;
    fold_partials_block(partials, nblk, f.cp, c, s, q, sh);
    if (lane != 0) return;
    if (f.dbeta) f.dbeta[c] = (float)s;
    if (f.dgamma) f.dgamma[c] = (float)q;
    if (f.sums) { f.sums[c] = s; f.sums[f.cp + c] = q; }
    f.c1[c] = f.eval_mode ? 0.f : (float)(s / (double)f.count);
    f.c2[c] = f.eval_mode ? 0.f : (float)(q / (double)f.count);
}

__global__ void bn_bwd_coeffs_kernel(const double* __restrict__ sums, int cp, int channels, const double* __restrict__ count_dev, float* c1, float* c2) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= cp) return;
    const double count = *count_dev;
    c1[c] = c < channels ? (float)(sums[c] / count) : 0.f;
    c2[c] = c < channels ? (float)(sums[cp + c] / count) : 0.f;
}

// g = dz * [relu mask];  partial sums of g and g*xhat.
template <typename T>
__global__ void __launch_bounds__(256, MSX_BN_MINBLOCKS) bn_bwd_reduce_kernel(const T* __restrict__ x, int ldx, const T* __restrict__ dz, int lddz,
                                                            int64_t pixels, int cp, int vch, int channels,
                                                            const float* __restrict__ mean, const float* __restrict__ invstd,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            int relu, double* partials, const BnBwdFinalize fin) {
    constexpr int VEC = Vec<T>::N;
    __shared__ double shm[2 * VEC * 256 / kLaneReduceRounds];
    __shared__ int last_flag;
    const ChanThread t = chan_thread<VEC>(cp, vch);
    float mu[VEC], is[VEC], ga[VEC], be[VEC];
    double acc[2 * VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        const int c = t.c0 + j;
        const bool ok = t.active && c < channels;
        mu[j] = ok ? mean[c] : 0.f;
        is[j] = ok ? invstd[c] : 0.f;
        ga[j] = ok ? gamma[c] : 0.f;
        be[j] = ok ? beta[c] : 0.f;
        acc[j] = 0.0; acc[VEC + j] = 0.0;
    }
    if (t.active) {
        const int64_t step = (int64_t)gridDim.x * t.lanes;
        int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane;
        float fs[VEC], fq[VEC];
        auto body = [&](const float (&v)[VEC], const float (&g)[VEC], bool first) {
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                const float xh = (v[j] - mu[j]) * is[j];
                const float out = (v[j] - mu[j]) * (is[j] * ga[j]) + be[j];
                const float gg = (relu && !(out > 0.f)) ? 0.f : g[j];
                fs[j] = first ? gg : fs[j] + gg;
                fq[j] = first ? gg * xh : fmaf(gg, xh, fq[j]);
            }
        };
        auto flush = [&]() {
#pragma unroll
            for (int j = 0; j < VEC; ++j) { acc[j] += (double)fs[j]; acc[VEC + j] += (double)fq[j]; }
        };
        for (; p + 3 * step < pixels; p += 4 * step) {           // 8 independent 16 B loads in flight per thread
            float v[4][VEC], g[4][VEC];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                Vec<T>::load(x + (p + u * step) * ldx + t.c0, v[u]);
                Vec<T>::load(dz + (p + u * step) * lddz + t.c0, g[u]);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) body(v[u], g[u], u == 0);
            flush();                                             // one fp32->fp64 conversion per 4 pixels
        }
        for (; p < pixels; p += step) {
            float v[VEC], g[VEC];
            Vec<T>::load(x + p * ldx + t.c0, v);
            Vec<T>::load(dz + p * lddz + t.c0, g);
            body(v, g, true);
            flush();
        }
    }
    lane_reduce<2 * VEC>(acc, shm, vch, t.lane, t.lanes);
    if (t.active && t.lane == 0) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            partials[((int64_t)blockIdx.x * 2 + 0) * cp + t.c0 + j] = acc[j];
            partials[((int64_t)blockIdx.x * 2 + 1) * cp + t.c0 + j] = acc[VEC + j];
        }
    }
    if (fin.ticket && ticket_is_last(fin.ticket, gridDim.x * gridDim.y, &last_flag)) bn_bwd_finalize_all(fin, partials, gridDim.x);
}

template <typename T>
__global__ void __launch_bounds__(256) bn_bwd_apply_kernel(const T* __restrict__ x, int ldx, const T* __restrict__ dz, int lddz,
                                                           T* __restrict__ dx, int lddx, int64_t pixels, int cp, int vch, int channels,
                                                           const float* __restrict__ mean, const float* __restrict__ invstd,
                                                           const float* __restrict__ gamma, const float* __restrict__ beta,
                                                           const float* __restrict__ c1, const float* __restrict__ c2, int relu) {
    constexpr int VEC = Vec<T>::N;
    const ChanThread t = chan_thread<VEC>(cp, vch);
    if (!t.active) return;
    float mu[VEC], is[VEC], ga[VEC], be[VEC], k1[VEC], k2[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        const int c = t.c0 + j;
        const bool ok = c < channels;
        mu[j] = ok ? mean[c] : 0.f;
        is[j] = ok ? invstd[c] : 0.f;
        ga[j] = ok ? gamma[c] : 0.f;
        be[j] = ok ? beta[c] : 0.f;
        k1[j] = ok ? c1[c] : 0.f;
        k2[j] = ok ? c2[c] : 0.f;
    }
    const int64_t step = (int64_t)gridDim.x * t.lanes;
    int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane;
    auto body = [&](float (&v)[VEC], const float (&g)[VEC]) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            const float xh = (v[j] - mu[j]) * is[j];
            const float out = (v[j] - mu[j]) * (is[j] * ga[j]) + be[j];
            const float gg = (relu && !(out > 0.f)) ? 0.f : g[j];
            v[j] = (ga[j] * is[j]) * (gg - k1[j] - xh * k2[j]);
        }
    };
    for (; p + step < pixels; p += 2 * step) {
        float v0[VEC], g0[VEC], v1[VEC], g1[VEC];
        Vec<T>::load(x + p * ldx + t.c0, v0);
        Vec<T>::load(dz + p * lddz + t.c0, g0);
        Vec<T>::load(x + (p + step) * ldx + t.c0, v1);
        Vec<T>::load(dz + (p + step) * lddz + t.c0, g1);
        body(v0, g0);
        body(v1, g1);
        Vec<T>::store(dx + p * lddx + t.c0, v0);
        Vec<T>::store(dx + (p + step) * lddx + t.c0, v1);
    }
    for (; p < pixels; p += step) {
        float v[VEC], g[VEC];
        Vec<T>::load(x + p * ldx + t.c0, v);
        Vec<T>::load(dz + p * lddz + t.c0, g);
        body(v, g);
        Vec<T>::store(dx + p * lddx + t.c0, v);
    }
}

// ---- bf16 fast paths of the two BatchNorm-backward passes -------------------------------------------------------------
// ncu (round 2, C = 17 at 16 x 352 x 352): the generic kernels above sat at 39 % (reduce) and 56 % (apply) of DRAM peak with
// 147 / 98 registers per thread - ONE or two CTAs per SM, 12-23 % of the warp slots, i.e. too few bytes in flight, not a
// bandwidth limit.  These variants keep the per-channel state small (fp32 partial sums for the few dozen pixels a thread
// visits, converted to double only for the cross-thread tree; folded coefficients), hold loaded vectors packed, and are
// compiled for three CTAs per SM with eight 16-byte loads in flight per thread.  fp32 tensors (the parity mode) keep the
// double-accumulating generic kernels.
__device__ __forceinline__ uint4 ldg16(const bf16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void unpack8(const uint4& r, float (&v)[8]) {
    v[0] = __uint_as_float(r.x << 16); v[1] = __uint_as_float(r.x & 0xffff0000u);
    v[2] = __uint_as_float(r.y << 16); v[3] = __uint_as_float(r.y & 0xffff0000u);
    v[4] = __uint_as_float(r.z << 16); v[5] = __uint_as_float(r.z & 0xffff0000u);
    v[6] = __uint_as_float(r.w << 16); v[7] = __uint_as_float(r.w & 0xffff0000u);
}

// 16 doubles through an 8-wide tree in two rounds (16 KB of shared memory instead of 32)
__device__ inline void lane_reduce_2x8(double (&v)[16], double* sh, int vch, int lane, int lanes) {
    int p2 = 1;
    while (p2 < lanes) p2 <<= 1;
#pragma unroll
    for (int r = 0; r < 2; ++r) {
        if (r) __syncthreads();
#pragma unroll
        for (int j = 0; j < 8; ++j) sh[j * 256 + threadIdx.x] = v[r * 8 + j];
        __syncthreads();
        for (int off = p2 >> 1; off > 0; off >>= 1) {
            if (lane < off && lane + off < lanes) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sh[j * 256 + threadIdx.x] += sh[j * 256 + threadIdx.x + off * vch];
            }
            __syncthreads();
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) v[r * 8 + j] = sh[j * 256 + threadIdx.x];
    }
}

// 16 floats per thread summed over the threads of the CTA that hold the same channel vector (threadIdx.x % vch, vch <= 32):
// inside a warp by shuffles between threads vch * 2^k apart, across the 8 warps through shared memory - two block barriers
// instead of the fourteen of the double-precision tree above, which was a quarter of the kernel at 88 x 88 (ncu, round 2).
// The per-thread fp32 sums cover a few dozen pixels; 85 lanes more in a pairwise fp32 tree add ~1e-6 relative error, and the
// fold over CTAs stays in double.  Result in the threads with threadIdx.x < vch.
__device__ inline void lane_reduce_f32(float (&v)[16], float* sh, int vch) {
    const int i = threadIdx.x & 31, w = threadIdx.x >> 5;
    int smax = vch;
    while (smax < 16) smax <<= 1;
    for (int s = smax; s >= vch; s >>= 1) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float o = __shfl_down_sync(0xffffffffu, v[j], s & 31);
            if (i + s < 32) v[j] += o;
        }
    }
    if (i < vch) {
        const int c = (32 * w + i) % vch;
#pragma unroll
        for (int j = 0; j < 16; ++j) sh[(w * 16 + j) * 32 + c] = v[j];
    }
    __syncthreads();
    if (threadIdx.x < vch) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            float t = 0.f;
#pragma unroll
            for (int ww = 0; ww < 8; ++ww) t += sh[(ww * 16 + j) * 32 + threadIdx.x];
            v[j] = t;
        }
    }
}

__global__ void __launch_bounds__(256, 3) bn_bwd_reduce_bf16_kernel(const bf16* __restrict__ x, int ldx, const bf16* __restrict__ dz, int lddz,
                                                                    int64_t pixels, int cp, int vch, int channels,
                                                                    const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                    const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                    int relu, double* partials, const BnBwdFinalize fin) {
    __shared__ double shm[8 * 256];
    __shared__ int last_flag;
    pdl_wait();
    pdl_trigger();
    const ChanThread t = chan_thread<8>(cp, vch);
    float mu[8], sc[8], be[8], fs[8], fq[8];          // sc = invstd * gamma; fq = sum g * (x - mean), scaled by invstd at the end
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = t.c0 + j;
        const bool ok = t.active && c < channels;
        mu[j] = ok ? mean[c] : 0.f;
        sc[j] = ok ? invstd[c] * gamma[c] : 0.f;
        be[j] = ok ? beta[c] : 0.f;
        fs[j] = 0.f; fq[j] = 0.f;
    }
    if (t.active) {
        const int64_t step = (int64_t)gridDim.x * t.lanes;
        int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane;
        auto body = [&](const uint4& rv, const uint4& rg) {
            float v[8], g[8];
            unpack8(rv, v);
            unpack8(rg, g);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const float d = v[j] - mu[j];
                const float out = d * sc[j] + be[j];             // the forward's expression: same ReLU mask
                const float gg = (relu && !(out > 0.f)) ? 0.f : g[j];
                fs[j] += gg;
                fq[j] = fmaf(gg, d, fq[j]);
            }
        };
        for (; p + 3 * step < pixels; p += 4 * step) {           // 8 independent 16 B loads in flight per thread
            uint4 rv[4], rg[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                rv[u] = ldg16(x + (p + u * step) * ldx + t.c0);
                rg[u] = ldg16(dz + (p + u * step) * lddz + t.c0);
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) body(rv[u], rg[u]);
        }
        for (; p < pixels; p += step) body(ldg16(x + p * ldx + t.c0), ldg16(dz + p * lddz + t.c0));
    }
    pdl_trigger_late();
    double acc[16];
    if (vch <= 32) {
        float v16[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) { v16[j] = fs[j]; v16[8 + j] = fq[j]; }
        lane_reduce_f32(v16, reinterpret_cast<float*>(shm), vch);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = t.c0 + j;
            acc[j] = (double)v16[j];
            acc[8 + j] = (t.active && c < channels) ? (double)v16[8 + j] * (double)invstd[c] : 0.0;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int c = t.c0 + j;
            acc[j] = (double)fs[j];
            acc[8 + j] = (t.active && c < channels) ? (double)fq[j] * (double)invstd[c] : 0.0;
        }
        lane_reduce_2x8(acc, shm, vch, t.lane, t.lanes);
    }
    if (t.active && t.lane == 0) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            partials[((int64_t)blockIdx.x * 2 + 0) * cp + t.c0 + j] = acc[j];
            partials[((int64_t)blockIdx.x * 2 + 1) * cp + t.c0 + j] = acc[8 + j];
        }
    }
    if (fin.ticket && ticket_is_last(fin.ticket, gridDim.x * gridDim.y, &last_flag)) bn_bwd_finalize_all(fin, partials, gridDim.x);
}

// dx = gamma*invstd*(g - c1 - xhat*c2) with the per-channel constants folded: dx = a*g + e*(x - mean) + f,
// a = gamma*invstd, e = -a*invstd*c2, f = -a*c1
__global__ void __launch_bounds__(256, 3) bn_bwd_apply_bf16_kernel(const bf16* __restrict__ x, int ldx, const bf16* __restrict__ dz, int lddz,
                                                                   bf16* __restrict__ dx, int lddx, int64_t pixels, int cp, int vch, int channels,
                                                                   const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                   const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                   const float* __restrict__ c1, const float* __restrict__ c2, int relu) {
    pdl_wait();
    pdl_trigger();
    const ChanThread t = chan_thread<8>(cp, vch);
    if (!t.active) return;
    float mu[8], sc[8], be[8], e[8], f[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = t.c0 + j;
        const bool ok = c < channels;
        const float is = ok ? invstd[c] : 0.f;
        mu[j] = ok ? mean[c] : 0.f;
        sc[j] = ok ? is * gamma[c] : 0.f;
        be[j] = ok ? beta[c] : 0.f;
        e[j] = ok ? -sc[j] * is * c2[c] : 0.f;
        f[j] = ok ? -sc[j] * c1[c] : 0.f;
    }
    const int64_t step = (int64_t)gridDim.x * t.lanes;
    int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane;
    auto body = [&](const uint4& rv, const uint4& rg, bf16* out) {
        float v[8], g[8];
        unpack8(rv, v);
        unpack8(rg, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float d = v[j] - mu[j];
            const float o = d * sc[j] + be[j];
            const float gg = (relu && !(o > 0.f)) ? 0.f : g[j];
            v[j] = fmaf(sc[j], gg, fmaf(e[j], d, f[j]));
        }
        Vec<bf16>::store(out, v);
    };
    for (; p + 3 * step < pixels; p += 4 * step) {
        uint4 rv[4], rg[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            rv[u] = ldg16(x + (p + u * step) * ldx + t.c0);
            rg[u] = ldg16(dz + (p + u * step) * lddz + t.c0);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) body(rv[u], rg[u], dx + (p + u * step) * lddx + t.c0);
    }
    for (; p < pixels; p += step) body(ldg16(x + p * ldx + t.c0), ldg16(dz + p * lddz + t.c0), dx + p * lddx + t.c0);
}

// y = relu((x - mean) * (invstd * gamma) + beta), bf16: packed loads, four pixels in flight, grid = one resident wave
__global__ void __launch_bounds__(256, 4) bn_apply_bf16_kernel(const bf16* __restrict__ x, int ldx, bf16* __restrict__ z, int ldz,
                                                               int64_t pixels, int cp, int vch, int channels,
                                                               const float* __restrict__ mean, const float* __restrict__ invstd,
                                                               const float* __restrict__ gamma, const float* __restrict__ beta, int relu) {
    pdl_wait();
    pdl_trigger();
    const ChanThread t = chan_thread<8>(cp, vch);
    if (!t.active) return;
    float mu[8], sc[8], sh[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const int c = t.c0 + j;
        const bool ok = c < channels;
        mu[j] = ok ? mean[c] : 0.f;
        sc[j] = ok ? invstd[c] * gamma[c] : 0.f;
        sh[j] = ok ? beta[c] : 0.f;
    }
    const int64_t step = (int64_t)gridDim.x * t.lanes;
    int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane;
    auto body = [&](const uint4& rv, bf16* out) {
        float v[8];
        unpack8(rv, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float r = (v[j] - mu[j]) * sc[j] + sh[j];
            v[j] = relu ? fmaxf(r, 0.f) : r;
        }
        Vec<bf16>::store(out, v);
    };
    for (; p + 3 * step < pixels; p += 4 * step) {
        uint4 rv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) rv[u] = ldg16(x + (p + u * step) * ldx + t.c0);
#pragma unroll
        for (int u = 0; u < 4; ++u) body(rv[u], z + (p + u * step) * ldz + t.c0);
    }
    for (; p < pixels; p += step) body(ldg16(x + p * ldx + t.c0), z + p * ldz + t.c0);
}

// ---- n-way sum (+ nearest 2x upsample of input 0) with fused statistics --------------------
struct SumArgs {
    const void* x[8];
    int ld[8];
    int n;
};
template <typename T, bool STATS>
__global__ void __launch_bounds__(256, MSX_BN_MINBLOCKS) sum_stats_kernel(SumArgs a, int up2x, int H, int W, T* __restrict__ s, int lds,
                                                        int64_t pixels, int cp, int vch, double* partials,
                                                        const float* __restrict__ scale = nullptr, const float* __restrict__ shift = nullptr,
                                                        int relu = 0, const BnFinalize fin = BnFinalize{}) {
    constexpr int VEC = Vec<T>::N;
    __shared__ double sh[STATS ? 2 * VEC * 256 / kLaneReduceRounds : 1];
    __shared__ int last_flag;
    const ChanThread t = chan_thread<VEC>(cp, vch);
    // inference: eval-mode BN affine (+ReLU) applied to the sum before it is stored (STATS == false only)
    float asc[VEC], ash[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
        asc[j] = (!STATS && scale && t.active) ? scale[t.c0 + j] : 1.f;
        ash[j] = (!STATS && scale && t.active) ? shift[t.c0 + j] : 0.f;
    }
    double acc[2 * VEC];
    float fs[VEC], fq[VEC];
    int pending = 0;
#pragma unroll
    for (int j = 0; j < 2 * VEC; ++j) acc[j] = 0.0;
#pragma unroll
    for (int j = 0; j < VEC; ++j) { fs[j] = 0.f; fq[j] = 0.f; }
    if (t.active) {
        for (int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane; p < pixels; p += (int64_t)gridDim.x * t.lanes) {
            int64_t p0 = p;
            if (up2x) {
                const int xw = (int)(p % W);
                const int64_t r = p / W;
                const int yh = (int)(r % H);
                const int64_t nn = r / H;
                p0 = (nn * (H >> 1) + (yh >> 1)) * (W >> 1) + (xw >> 1);
            }
            float v[VEC];
            Vec<T>::load(static_cast<const T*>(a.x[0]) + p0 * a.ld[0] + t.c0, v);
            // inputs in groups of four: four independent 16 B loads in flight, then the adds
#pragma unroll
            for (int i0 = 1; i0 < 8; i0 += 4) {
                if (i0 < a.n) {
                    float u[4][VEC];
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (i0 + i < a.n) Vec<T>::load(static_cast<const T*>(a.x[i0 + i]) + p * a.ld[i0 + i] + t.c0, u[i]);
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        if (i0 + i < a.n) {
#pragma unroll
                            for (int j = 0; j < VEC; ++j) v[j] += u[i][j];
                        }
                }
            }
            if (!STATS && (scale || relu)) {
#pragma unroll
                for (int j = 0; j < VEC; ++j) {
                    if (scale) v[j] = fmaf(v[j], asc[j], ash[j]);
                    if (relu) v[j] = fmaxf(v[j], 0.f);
                }
            }
            Vec<T>::store(s + p * lds + t.c0, v);
            if (!STATS) continue;
            // statistics of the value as stored (bf16 rounding included), so that apply/backward,
            // which re-read s, see the same numbers; fp32 over 4 pixels, then one conversion to double
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                const float w = to_float(from_float<T>(v[j]));
                fs[j] += w;
                fq[j] = fmaf(w, w, fq[j]);
            }
            if (++pending == 4) {
#pragma unroll
                for (int j = 0; j < VEC; ++j) { acc[j] += (double)fs[j]; acc[VEC + j] += (double)fq[j]; fs[j] = 0.f; fq[j] = 0.f; }
                pending = 0;
            }
        }
#pragma unroll
        for (int j = 0; j < VEC; ++j) { acc[j] += (double)fs[j]; acc[VEC + j] += (double)fq[j]; }
    }
    if (!STATS) return;
    lane_reduce<2 * VEC>(acc, sh, vch, t.lane, t.lanes);
    if (t.active && t.lane == 0 && partials) {
#pragma unroll
        for (int j = 0; j < VEC; ++j) {
            partials[((int64_t)blockIdx.x * 2 + 0) * cp + t.c0 + j] = acc[j];
            partials[((int64_t)blockIdx.x * 2 + 1) * cp + t.c0 + j] = acc[VEC + j];
        }
    }
    if (partials && fin.ticket && ticket_is_last(fin.ticket, gridDim.x * gridDim.y, &last_flag)) bn_finalize_all(fin, gridDim.x);
}

// bf16 fast path of the n-way sum (see the note above the BatchNorm-backward fast paths): U pixels x NMAX inputs = eight
// packed 16-byte loads in flight per thread, fp32 running statistics of the stored (bf16-rounded) sums.
template <int U, int NMAX, bool STATS>
__device__ __forceinline__ void sum_pixels_bf16(const SumArgs& a, int up2x, int H, int W, const int64_t (&p)[U], int c0, bf16* __restrict__ s, int lds,
                                                const float (&asc)[8], const float (&ash)[8], bool affine, int relu, float (&fs)[8], float (&fq)[8]) {
    uint4 r[U][NMAX];
#pragma unroll
    for (int u = 0; u < U; ++u) {
        int64_t p0 = p[u];
        if (up2x) {
            const int xw = (int)(p[u] % W);
            const int64_t q = p[u] / W;
            const int yh = (int)(q % H);
            p0 = ((q / H) * (H >> 1) + (yh >> 1)) * (W >> 1) + (xw >> 1);
        }
        r[u][0] = ldg16(static_cast<const bf16*>(a.x[0]) + p0 * a.ld[0] + c0);
#pragma unroll
        for (int i = 1; i < NMAX; ++i)
            if (i < a.n) r[u][i] = ldg16(static_cast<const bf16*>(a.x[i]) + p[u] * a.ld[i] + c0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
        float v[8];
        unpack8(r[u][0], v);
#pragma unroll
        for (int i = 1; i < NMAX; ++i)
            if (i < a.n) {
                float w[8];
                unpack8(r[u][i], w);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j] += w[j];
            }
        if (!STATS && (affine || relu)) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (affine) v[j] = fmaf(v[j], asc[j], ash[j]);
                if (relu) v[j] = fmaxf(v[j], 0.f);
            }
        }
        uint32_t w4[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
            w4[j] = *reinterpret_cast<uint32_t*>(&h);
        }
        *reinterpret_cast<uint4*>(s + p[u] * lds + c0) = make_uint4(w4[0], w4[1], w4[2], w4[3]);
        if (STATS) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float lo = __uint_as_float(w4[j] << 16), hi = __uint_as_float(w4[j] & 0xffff0000u);
                fs[2 * j] += lo; fq[2 * j] = fmaf(lo, lo, fq[2 * j]);
                fs[2 * j + 1] += hi; fq[2 * j + 1] = fmaf(hi, hi, fq[2 * j + 1]);
            }
        }
    }
}

template <bool STATS>
__global__ void __launch_bounds__(256, 3) sum_bf16_kernel(SumArgs a, int up2x, int H, int W, bf16* __restrict__ s, int lds, int64_t pixels,
                                                          int cp, int vch, double* partials, const float* __restrict__ scale,
                                                          const float* __restrict__ shift, int relu, const BnFinalize fin) {
    __shared__ double sh[STATS ? 8 * 256 : 1];
    __shared__ int last_flag;
    pdl_wait();
    pdl_trigger();
    const ChanThread t = chan_thread<8>(cp, vch);
    float asc[8], ash[8], fs[8], fq[8];
    const bool affine = !STATS && scale != nullptr;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        asc[j] = (affine && t.active) ? scale[t.c0 + j] : 1.f;
        ash[j] = (affine && t.active) ? shift[t.c0 + j] : 0.f;
        fs[j] = 0.f; fq[j] = 0.f;
    }
    if (t.active) {
        const int64_t step = (int64_t)gridDim.x * t.lanes;
        int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane;
        if (a.n <= 2) {
            for (; p + 3 * step < pixels; p += 4 * step) {
                const int64_t pp[4] = {p, p + step, p + 2 * step, p + 3 * step};
                sum_pixels_bf16<4, 2, STATS>(a, up2x, H, W, pp, t.c0, s, lds, asc, ash, affine, relu, fs, fq);
            }
        } else if (a.n <= 4) {
            for (; p + step < pixels; p += 2 * step) {
                const int64_t pp[2] = {p, p + step};
                sum_pixels_bf16<2, 4, STATS>(a, up2x, H, W, pp, t.c0, s, lds, asc, ash, affine, relu, fs, fq);
            }
        }
        for (; p < pixels; p += step) {
            const int64_t pp[1] = {p};
            sum_pixels_bf16<1, 8, STATS>(a, up2x, H, W, pp, t.c0, s, lds, asc, ash, affine, relu, fs, fq);
        }
    }
    if (!STATS) return;
    pdl_trigger_late();
    double acc[16];
    if (vch <= 32) {
        float v16[16];
#pragma unroll
        for (int j = 0; j < 8; ++j) { v16[j] = fs[j]; v16[8 + j] = fq[j]; }
        lane_reduce_f32(v16, reinterpret_cast<float*>(sh), vch);
#pragma unroll
        for (int j = 0; j < 16; ++j) acc[j] = (double)v16[j];
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc[j] = (double)fs[j]; acc[8 + j] = (double)fq[j]; }
        lane_reduce_2x8(acc, sh, vch, t.lane, t.lanes);
    }
    if (t.active && t.lane == 0 && partials) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            partials[((int64_t)blockIdx.x * 2 + 0) * cp + t.c0 + j] = acc[j];
            partials[((int64_t)blockIdx.x * 2 + 1) * cp + t.c0 + j] = acc[8 + j];
        }
    }
    if (partials && fin.ticket && ticket_is_last(fin.ticket, gridDim.x * gridDim.y, &last_flag)) bn_finalize_all(fin, gridDim.x);
}

template <typename T>
__global__ void __launch_bounds__(256) upsample2x_bwd_kernel(const T* __restrict__ ds, int ldds, int H, int W,
                                                             T* __restrict__ da, int ldda, int64_t pixels_lo, int cp, int vch) {
    constexpr int VEC = Vec<T>::N;
    const ChanThread t = chan_thread<VEC>(cp, vch);
    if (!t.active) return;
    const int h2 = H >> 1, w2 = W >> 1;
    for (int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane; p < pixels_lo; p += (int64_t)gridDim.x * t.lanes) {
        const int xw = (int)(p % w2);
        const int64_t r = p / w2;
        const int yh = (int)(r % h2);
        const int64_t nn = r / h2;
        const int64_t base = (nn * H + 2 * yh) * W + 2 * xw;
        float v[VEC], u[VEC];
        Vec<T>::load(ds + base * ldds + t.c0, v);
        Vec<T>::load(ds + (base + 1) * ldds + t.c0, u);
#pragma unroll
        for (int j = 0; j < VEC; ++j) v[j] += u[j];
        Vec<T>::load(ds + (base + W) * ldds + t.c0, u);
#pragma unroll
        for (int j = 0; j < VEC; ++j) v[j] += u[j];
        Vec<T>::load(ds + (base + W + 1) * ldds + t.c0, u);
#pragma unroll
        for (int j = 0; j < VEC; ++j) v[j] += u[j];
        Vec<T>::store(da + p * ldda + t.c0, v);
    }
}

// ---- MaxPool2d(3, 2, 1) -------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const T* __restrict__ x, int ldx, int H, int W, int OH, int OW,
                                                          T* __restrict__ z, int ldz, int8_t* __restrict__ idx, int ldidx,
                                                          int64_t opix, int cp, int vch) {
    constexpr int VEC = Vec<T>::N;
    const ChanThread t = chan_thread<VEC>(cp, vch);
    if (!t.active) return;
    for (int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane; p < opix; p += (int64_t)gridDim.x * t.lanes) {
        const int ox = (int)(p % OW);
        const int64_t r = p / OW;
        const int oy = (int)(r % OH);
        const int64_t nn = r / OH;
        float best[VEC];
        int bi[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) { best[j] = -INFINITY; bi[j] = -1; }
#pragma unroll
        for (int ky = 0; ky < 3; ++ky) {
            const int iy = 2 * oy - 1 + ky;
            if (iy < 0 || iy >= H) continue;
#pragma unroll
            for (int kx = 0; kx < 3; ++kx) {
                const int ix = 2 * ox - 1 + kx;
                if (ix < 0 || ix >= W) continue;
                float v[VEC];
                Vec<T>::load(x + ((nn * H + iy) * W + ix) * ldx + t.c0, v);
#pragma unroll
                for (int j = 0; j < VEC; ++j)
                    if (bi[j] < 0 || v[j] > best[j]) { best[j] = v[j]; bi[j] = ky * 3 + kx; }
            }
        }
        Vec<T>::store(z + p * ldz + t.c0, best);
        int8_t* ip = idx + p * ldidx + t.c0;
#pragma unroll
        for (int j = 0; j < VEC; ++j) ip[j] = (int8_t)bi[j];
    }
}

template <typename T>
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const T* __restrict__ dz, int lddz, const int8_t* __restrict__ idx, int ldidx,
                                                          int H, int W, int OH, int OW, T* __restrict__ dx, int lddx,
                                                          int64_t ipix, int cp, int vch) {
    constexpr int VEC = Vec<T>::N;
    const ChanThread t = chan_thread<VEC>(cp, vch);
    if (!t.active) return;
    for (int64_t p = (int64_t)blockIdx.x * t.lanes + t.lane; p < ipix; p += (int64_t)gridDim.x * t.lanes) {
        const int ix = (int)(p % W);
        const int64_t r = p / W;
        const int iy = (int)(r % H);
        const int64_t nn = r / H;
        float acc[VEC];
#pragma unroll
        for (int j = 0; j < VEC; ++j) acc[j] = 0.f;
        const int oy0 = iy >> 1, oy1 = (iy + 1) >> 1;     // windows 2*oy-1 .. 2*oy+1 containing iy
        const int ox0 = ix >> 1, ox1 = (ix + 1) >> 1;
        for (int oy = oy0; oy <= oy1; ++oy) {
            if (oy >= OH) continue;
            const int ky = iy - (2 * oy - 1);
            for (int ox = ox0; ox <= ox1; ++ox) {
                if (ox >= OW) continue;
                const int pos = ky * 3 + (ix - (2 * ox - 1));
                const int64_t op = (nn * OH + oy) * OW + ox;
                float g[VEC];
                Vec<T>::load(dz + op * lddz + t.c0, g);
                const int8_t* ip = idx + op * ldidx + t.c0;
#pragma unroll
                for (int j = 0; j < VEC; ++j)
                    if (ip[j] == pos) acc[j] += g[j];
            }
        }
        Vec<T>::store(dx + p * lddx + t.c0, acc);
    }
}

// ---- layout boundary ----------------------------------------------------------------------
template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ src, int C, int64_t hw, int64_t pixels, T* __restrict__ dst, int ld, int cp) {
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < pixels; p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t nn = p / hw, q = p % hw;
        for (int c0 = 0; c0 < cp; c0 += 4) {
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = (c0 + j < C) ? src[(nn * C + c0 + j) * hw + q] : 0.f;
            store4(dst + p * ld + c0, v);
        }
    }
}
template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ src, int ld, int C, int64_t hw, int64_t pixels, float* __restrict__ dst) {
    for (int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; p < pixels; p += (int64_t)gridDim.x * blockDim.x) {
        const int64_t nn = p / hw, q = p % hw;
        for (int c = 0; c < C; ++c) dst[(nn * C + c) * hw + q] = to_float(src[p * ld + c]);
    }
}

// per-channel plain sum -> float (ConvTranspose bias gradient)
__global__ void channel_sum_finalize_kernel(const double* __restrict__ partials, int nblk, int cp, int channels, float* out) {
    const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (c >= channels) return;
    double s, q;
    fold_partials(partials, nblk, cp, c, s, q);
    if ((threadIdx.x & 31) == 0) out[c] = (float)s;
}

// CTAs of `kernel` (256 threads, no dynamic smem) that one resident wave holds on the current device; cached per kernel
inline int resident_ctas(const void* kernel) {
    static std::mutex mu;
    static std::map<std::pair<int, const void*>, int> cache;
    int dev = 0;
    cudaGetDevice(&dev);
    std::lock_guard<std::mutex> lk(mu);
    auto it = cache.find({dev, kernel});
    if (it != cache.end()) return it->second;
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 256, 0) != cudaSuccess || per_sm < 1) per_sm = 2;
    const int n = per_sm * kNumSMs;
    cache[{dev, kernel}] = n;
    return n;
}
inline int wave_grid(const void* kernel, int64_t pixels, int lanes, int per_iter, int cap = 1 << 30) {
    int64_t g = cdiv64(pixels, (int64_t)lanes * per_iter);
    const int r = resident_ctas(kernel);
    if (g > r) g = r;
    if (g > cap) g = cap;
    return g < 1 ? 1 : (int)g;
}

inline int check_ld(int ld, int channels, const char* what) {
    if (ld % 8 != 0 || ld < round_up(channels, 8))
        return set_error(MSX_ERR_INVALID, "%s: pixel stride %d must be a multiple of 8 and >= round_up(%d, 8)", what, ld, channels);
    return MSX_OK;
}

}  // namespace

namespace {
template <bool STATS>
int sum_launch_bf16(const SumArgs& a, int up2x, int h, int w, void* s, int lds, int64_t pixels, int channels, double* partials,
                    const float* scale, const float* shift, int relu, const BnFinalize& fin, int* nblk_out, cudaStream_t st) {
    const ChanCfg c = chan_cfg<8>(channels);
    const int per_iter = a.n <= 2 ? 4 : (a.n <= 4 ? 2 : 1);
    const int g = wave_grid(reinterpret_cast<const void*>(sum_bf16_kernel<STATS>), pixels, c.lanes, per_iter * 2, STATS ? kMaxPartialBlocks : (1 << 30));
    { ProfScope ps__(K_SUM, st, (double)pixels * channels * 2.0 * (a.n + 1.0), 0);
      launch_pdl(sum_bf16_kernel<STATS>, dim3(dim3(g, c.nchunk)), dim3(256), 0, st, a, up2x, h, w, static_cast<bf16*>(s), lds, pixels, c.cp, c.vch, partials, scale, shift, relu, fin); }
    if (nblk_out) *nblk_out = g;
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}
}  // namespace

// ---- launchers (internal) -------------------------------------------------------------------
size_t bn_partials_bytes(int channels) {
    const int cp = round_up(channels, 8);
    return (size_t)kMaxPartialBlocks * 2 * cp * sizeof(double) + 2 * (size_t)cp * sizeof(float) + 256;
}

int ticket_max_cp() {
    static const int v = [] { const char* e = getenv("MSX_TICKET_MAX_CP"); return e ? atoi(e) : 0; }();
    return v;
}

// fold nblk partial rows and finalize, as its own launch (what every reduction uses when its BnFinalize has no ticket)
int bn_finalize_launch(const BnFinalize& f, int nblk, cudaStream_t st) {
    { ProfScope ps__(K_BN_FINALIZE, st, 0, 0); launch_pdl(bn_finalize_kernel, dim3(f.cp), dim3(128), 0, st, nblk, f); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int bn_finalize_args(BnFinalize* f, double* partials, int channels, int64_t count, float eps, float momentum, float* mean,
                     float* invstd, float* running_mean, float* running_var, int64_t* nbt, double* sums) {
    f->partials = partials;
    f->ticket = nullptr;
    if (round_up(channels, 8) <= ticket_max_cp()) {
        f->ticket = next_ticket();
        if (!f->ticket) return set_error(MSX_ERR_CUDA, "reduction ticket array unavailable (cudaGetSymbolAddress failed)");
    }
    f->sums = sums;
    f->mean = mean; f->invstd = invstd; f->running_mean = running_mean; f->running_var = running_var;
    f->nbt = reinterpret_cast<long long*>(nbt);
    f->count = count;
    f->channels = channels; f->cp = round_up(channels, 8);
    f->eps = eps; f->momentum = momentum;
    return MSX_OK;
}

// fin == nullptr: per-CTA partial rows only (nblk_out rows); otherwise the launch's last CTA folds and finalizes them
int bn_stats_launch(int dtype, const void* x, int ld, int64_t pixels, int channels, double* partials, const BnFinalize* fin,
                    int* nblk_out, cudaStream_t st) {
    MSX_TRY(check_ld(ld, channels, "bn_stats"));
    BnFinalize f = fin ? *fin : BnFinalize{};
    MSX_DISPATCH_DTYPE(dtype, T, {
        const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
        const int g = reduce_grid(pixels, c.lanes);
        { ProfScope ps__(K_BN_STATS, st, (double)pixels * channels * (double)sizeof(T), 0); bn_stats_kernel<T><<<dim3(g, c.nchunk), 256, 0, st>>>(static_cast<const T*>(x), ld, pixels, c.cp, c.vch, partials, f); }
        if (nblk_out) *nblk_out = g;
        if (fin && !fin->ticket) MSX_TRY(bn_finalize_launch(f, g, st));
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int bn_finalize_sums_launch(const double* sums, const BnFinalize& f, cudaStream_t st) {
    { ProfScope ps__(K_BN_FINALIZE, st, 0, 0); bn_finalize_sums_kernel<<<(f.cp + 127) / 128, 128, 0, st>>>(sums, f); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int bn_eval_stats_launch(int channels, const float* rm, const float* rv, float eps, float* mean, float* invstd,
                         cudaStream_t st) {
    const int cp = round_up(channels, 8);
    { ProfScope ps__(K_BN_FINALIZE, st, 0, 0); bn_eval_stats_kernel<<<(cp + 127) / 128, 128, 0, st>>>(cp, channels, rm, rv, eps, mean, invstd); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int bn_eval_affine_launch(int channels, const float* rm, const float* rv, float eps, const float* gamma, const float* beta,
                          const float* conv_bias, float* mean, float* invstd, float* scale, float* shift, cudaStream_t st) {
    const int cp = round_up(channels, 8);
    { ProfScope ps__(K_BN_FINALIZE, st, 0, 0); bn_eval_affine_kernel<<<(cp + 127) / 128, 128, 0, st>>>(cp, channels, rm, rv, eps, gamma, beta, conv_bias, mean, invstd, scale, shift); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int bn_apply_launch(int dtype, const void* x, int ldx, void* z, int ldz, int64_t pixels, int channels,
                    const float* mean, const float* invstd, const float* gamma, const float* beta, int relu,
                    cudaStream_t st) {
    MSX_TRY(check_ld(ldx, channels, "bn_apply x"));
    MSX_TRY(check_ld(ldz, channels, "bn_apply z"));
    if (dtype == MSX_BF16) {
        const ChanCfg c = chan_cfg<8>(channels);
        const int g = wave_grid(reinterpret_cast<const void*>(bn_apply_bf16_kernel), pixels, c.lanes, 4);
        { ProfScope ps__(K_BN_APPLY, st, 2.0 * (double)pixels * channels * 2.0, 0); launch_pdl(bn_apply_bf16_kernel, dim3(dim3(g, c.nchunk)), dim3(256), 0, st, 
            static_cast<const bf16*>(x), ldx, static_cast<bf16*>(z), ldz, pixels, c.cp, c.vch, channels, mean, invstd, gamma, beta, relu); }
        MSX_LAUNCH_CHECK();
        return MSX_OK;
    }
    MSX_DISPATCH_DTYPE(dtype, T, {
        const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
        { ProfScope ps__(K_BN_APPLY, st, 2.0 * (double)pixels * channels * (double)sizeof(T), 0); bn_apply_kernel<T><<<dim3(stream_grid(pixels, c.lanes), c.nchunk), 256, 0, st>>>(
            static_cast<const T*>(x), ldx, static_cast<T*>(z), ldz, pixels, c.cp, c.vch, channels, mean, invstd, gamma, beta, relu); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

// The two halves of the BatchNorm(+ReLU) backward.  reduce: dgamma / dbeta and the coefficients c1, c2 (fp32[cp] each, in
// the caller's scratch) by the launch's last CTA; sums (nullable, [2][cp] doubles): the raw sums for SyncBatchNorm.
int bn_bwd_reduce_launch(int dtype, const void* x, int ldx, const void* dz, int lddz, int64_t pixels, int channels,
                         const float* gamma, const float* beta, const float* mean, const float* invstd, int flags,
                         float* dgamma, float* dbeta, double* partials, float* c1, float* c2, double* sums, cudaStream_t st) {
    const int relu = flags & MSX_BN_RELU;
    MSX_TRY(check_ld(ldx, channels, "bn_bwd x"));
    MSX_TRY(check_ld(lddz, channels, "bn_bwd dz"));
    BnBwdFinalize f;
    f.ticket = nullptr;
    if (round_up(channels, 8) <= ticket_max_cp()) {
        f.ticket = next_ticket();
        if (!f.ticket) return set_error(MSX_ERR_CUDA, "reduction ticket array unavailable (cudaGetSymbolAddress failed)");
    }
    f.sums = sums; f.dgamma = dgamma; f.dbeta = dbeta; f.c1 = c1; f.c2 = c2;
    f.count = pixels; f.channels = channels; f.cp = round_up(channels, 8); f.eval_mode = (flags & MSX_BN_EVAL) ? 1 : 0;
    if (dtype == MSX_BF16) {
        const ChanCfg c = chan_cfg<8>(channels);
        const int g = wave_grid(reinterpret_cast<const void*>(bn_bwd_reduce_bf16_kernel), pixels, c.lanes, 8, kMaxPartialBlocks);
        { ProfScope ps__(K_BN_BWD_REDUCE, st, 2.0 * (double)pixels * channels * 2.0, 0); launch_pdl(bn_bwd_reduce_bf16_kernel, dim3(dim3(g, c.nchunk)), dim3(256), 0, st, 
            static_cast<const bf16*>(x), ldx, static_cast<const bf16*>(dz), lddz, pixels, c.cp, c.vch, channels, mean, invstd, gamma, beta, relu, partials, f); }
        if (!f.ticket) { ProfScope ps__(K_BN_FINALIZE, st, 0, 0); launch_pdl(bn_bwd_finalize_kernel, dim3(f.cp), dim3(128), 0, st, partials, g, f); }
        MSX_LAUNCH_CHECK();
        return MSX_OK;
    }
    MSX_DISPATCH_DTYPE(dtype, T, {
        const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
        const int g = reduce_grid(pixels, c.lanes);
        { ProfScope ps__(K_BN_BWD_REDUCE, st, 2.0 * (double)pixels * channels * (double)sizeof(T), 0); bn_bwd_reduce_kernel<T><<<dim3(g, c.nchunk), 256, 0, st>>>(static_cast<const T*>(x), ldx, static_cast<const T*>(dz), lddz,
                                                                 pixels, c.cp, c.vch, channels, mean, invstd, gamma, beta, relu, partials, f); }
        if (!f.ticket) { ProfScope ps__(K_BN_FINALIZE, st, 0, 0); launch_pdl(bn_bwd_finalize_kernel, dim3(f.cp), dim3(128), 0, st, partials, g, f); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int bn_bwd_apply_launch(int dtype, const void* x, int ldx, const void* dz, int lddz, int64_t pixels, int channels,
                        const float* gamma, const float* beta, const float* mean, const float* invstd, int flags,
                        const float* c1, const float* c2, void* dx, int lddx, cudaStream_t st) {
    const int relu = flags & MSX_BN_RELU;
    MSX_TRY(check_ld(lddx, channels, "bn_bwd dx"));
    if (dtype == MSX_BF16) {
        const ChanCfg c = chan_cfg<8>(channels);
        const int g = wave_grid(reinterpret_cast<const void*>(bn_bwd_apply_bf16_kernel), pixels, c.lanes, 4);
        { ProfScope ps__(K_BN_BWD_APPLY, st, 3.0 * (double)pixels * channels * 2.0, 0); launch_pdl(bn_bwd_apply_bf16_kernel, dim3(dim3(g, c.nchunk)), dim3(256), 0, st, 
            static_cast<const bf16*>(x), ldx, static_cast<const bf16*>(dz), lddz, static_cast<bf16*>(dx), lddx, pixels, c.cp, c.vch, channels,
            mean, invstd, gamma, beta, c1, c2, relu); }
        MSX_LAUNCH_CHECK();
        return MSX_OK;
    }
    MSX_DISPATCH_DTYPE(dtype, T, {
        const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
        { ProfScope ps__(K_BN_BWD_APPLY, st, 3.0 * (double)pixels * channels * (double)sizeof(T), 0); bn_bwd_apply_kernel<T><<<dim3(stream_grid(pixels, c.lanes), c.nchunk), 256, 0, st>>>(
            static_cast<const T*>(x), ldx, static_cast<const T*>(dz), lddz, static_cast<T*>(dx), lddx, pixels, c.cp, c.vch, channels,
            mean, invstd, gamma, beta, c1, c2, relu); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int bn_bwd_launch(int dtype, const void* x, int ldx, const void* dz, int lddz, int64_t pixels, int channels,
                  const float* gamma, const float* beta, const float* mean, const float* invstd, int flags,
                  void* dx, int lddx, float* dgamma, float* dbeta, void* ws, size_t ws_bytes, cudaStream_t st) {
    if (ws_bytes < bn_partials_bytes(channels)) return set_error(MSX_ERR_WORKSPACE, "bn_bwd: workspace %zu < %zu", ws_bytes, bn_partials_bytes(channels));
    const int cp = round_up(channels, 8);
    double* partials = static_cast<double*>(ws);
    float* c1 = reinterpret_cast<float*>(partials + (size_t)kMaxPartialBlocks * 2 * cp);
    float* c2 = c1 + cp;
    MSX_TRY(bn_bwd_reduce_launch(dtype, x, ldx, dz, lddz, pixels, channels, gamma, beta, mean, invstd, flags, dgamma, dbeta, partials, c1, c2, nullptr, st));
    return bn_bwd_apply_launch(dtype, x, ldx, dz, lddz, pixels, channels, gamma, beta, mean, invstd, flags, c1, c2, dx, lddx, st);
}

int channel_sum_launch(int dtype, const void* x, int ld, int64_t pixels, int channels, float* out, void* ws,
                       size_t ws_bytes, cudaStream_t st) {
    if (ws_bytes < bn_partials_bytes(channels)) return set_error(MSX_ERR_WORKSPACE, "channel_sum: workspace too small");
    int nblk = 0;
    MSX_TRY(bn_stats_launch(dtype, x, ld, pixels, channels, static_cast<double*>(ws), nullptr, &nblk, st));
    { ProfScope ps__(K_MISC, st, 0, 0); channel_sum_finalize_kernel<<<(channels * 32 + 127) / 128, 128, 0, st>>>(static_cast<double*>(ws), nblk, round_up(channels, 8), channels, out); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

}  // namespace msx

using namespace msx;

// ============================================================================================
extern "C" {

size_t msx_ws_bytes_bn(int32_t channels) { return bn_partials_bytes(channels); }

int msx_sum_bn_act_fwd(int dtype, int n_in, const void* const* host_xs, const int32_t* host_ldxs, int up2x_first,
                       int32_t n, int32_t h, int32_t w, int32_t channels, void* s, int32_t lds,
                       const float* gamma, const float* beta, float* running_mean, float* running_var,
                       int64_t* num_batches_tracked, float eps, float momentum, int training, int relu,
                       void* z, int32_t ldz, float* save_mean, float* save_invstd, void* ws, size_t ws_bytes, void* stream) {
    cudaStream_t st = as_stream(stream);
    MSX_REQUIRE(n_in >= 1 && n_in <= 8, "sum_bn_act: n_in=%d not in [1,8]", n_in);
    MSX_REQUIRE(host_xs && host_ldxs && gamma && beta && z && save_mean && save_invstd, "sum_bn_act: null pointer");
    MSX_REQUIRE(!up2x_first || (h % 2 == 0 && w % 2 == 0), "sum_bn_act: nearest-2x needs even h, w (got %dx%d)", h, w);
    if (ws_bytes < bn_partials_bytes(channels)) return set_error(MSX_ERR_WORKSPACE, "sum_bn_act: workspace %zu < %zu", ws_bytes, bn_partials_bytes(channels));
    const int64_t pixels = (int64_t)n * h * w;
    double* partials = static_cast<double*>(ws);
    const void* src = host_xs[0];
    int ldsrc = host_ldxs[0];
    int nblk = 0;
    const bool need_sum = n_in > 1 || up2x_first;
    BnFinalize fin{};                 // training: the statistics kernel's last CTA finalizes (mean, invstd, running stats)
    if (training)
        MSX_TRY(bn_finalize_args(&fin, partials, channels, pixels, eps, momentum, save_mean, save_invstd, running_mean, running_var,
                                 num_batches_tracked, nullptr));
    if (need_sum && !training && s == nullptr) {
        // inference: one kernel sums the inputs, applies the eval-mode BN affine and the ReLU and writes z
        MSX_REQUIRE(running_mean && running_var, "sum_bn_act: eval mode needs running statistics");
        MSX_TRY(check_ld(ldz, channels, "sum_bn_act z"));
        float* scale = reinterpret_cast<float*>(ws);
        float* shift = scale + round_up(channels, 8);
        MSX_TRY(bn_eval_affine_launch(channels, running_mean, running_var, eps, gamma, beta, nullptr, save_mean, save_invstd, scale, shift, st));
        SumArgs a;
        a.n = n_in;
        for (int i = 0; i < n_in; ++i) {
            MSX_TRY(check_ld(host_ldxs[i], channels, "sum_bn_act x"));
            a.x[i] = host_xs[i];
            a.ld[i] = host_ldxs[i];
        }
        if (dtype == MSX_BF16) return sum_launch_bf16<false>(a, up2x_first, h, w, z, ldz, pixels, channels, nullptr, scale, shift, relu, BnFinalize{}, nullptr, st);
        MSX_DISPATCH_DTYPE(dtype, T, {
            const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
            { ProfScope ps__(K_SUM, st, (double)pixels * channels * (double)sizeof(T) * (n_in + 1.0), 0); sum_stats_kernel<T, false><<<dim3(stream_grid(pixels, c.lanes), c.nchunk), 256, 0, st>>>(
                a, up2x_first, h, w, static_cast<T*>(z), ldz, pixels, c.cp, c.vch, nullptr, scale, shift, relu); }
        });
        MSX_LAUNCH_CHECK();
        return MSX_OK;
    }
    if (need_sum) {
        MSX_REQUIRE(s != nullptr, "sum_bn_act: sum buffer required for n_in>1 or upsampling (training, or a backward may follow)");
        MSX_TRY(check_ld(lds, channels, "sum_bn_act s"));
        SumArgs a;
        a.n = n_in;
        for (int i = 0; i < n_in; ++i) {
            MSX_TRY(check_ld(host_ldxs[i], channels, "sum_bn_act x"));
            a.x[i] = host_xs[i];
            a.ld[i] = host_ldxs[i];
        }
        if (dtype == MSX_BF16) {
            MSX_TRY(sum_launch_bf16<true>(a, up2x_first, h, w, s, lds, pixels, channels, training ? partials : nullptr, nullptr, nullptr, 0, fin, &nblk, st));
        } else
        MSX_DISPATCH_DTYPE(dtype, T, {
            const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
            nblk = reduce_grid(pixels, c.lanes);
            { ProfScope ps__(K_SUM, st, (double)pixels * channels * (double)sizeof(T) * (n_in + 1.0), 0); sum_stats_kernel<T, true><<<dim3(nblk, c.nchunk), 256, 0, st>>>(a, up2x_first, h, w, static_cast<T*>(s), lds, pixels, c.cp, c.vch,
                                                                     training ? partials : nullptr, nullptr, nullptr, 0, fin); }
        });
        MSX_LAUNCH_CHECK();
        if (training && !fin.ticket) MSX_TRY(bn_finalize_launch(fin, nblk, st));
        src = s;
        ldsrc = lds;
    } else if (training) {
        MSX_TRY(bn_stats_launch(dtype, src, ldsrc, pixels, channels, partials, &fin, &nblk, st));
    }
    if (!training)
        MSX_TRY(bn_eval_stats_launch(channels, running_mean, running_var, eps, save_mean, save_invstd, st));
    return bn_apply_launch(dtype, src, ldsrc, z, ldz, pixels, channels, save_mean, save_invstd, gamma, beta, relu, st);
}

int msx_sum_n(int dtype, int n_in, const void* const* host_xs, const int32_t* host_ldxs, int64_t pixels, int32_t channels,
              void* out, int32_t ldout, void* stream) {
    MSX_REQUIRE(n_in >= 1 && n_in <= 8 && host_xs && host_ldxs && out, "sum_n: bad arguments");
    MSX_TRY(check_ld(ldout, channels, "sum_n out"));
    SumArgs a;
    a.n = n_in;
    for (int i = 0; i < n_in; ++i) {
        MSX_TRY(check_ld(host_ldxs[i], channels, "sum_n x"));
        a.x[i] = host_xs[i];
        a.ld[i] = host_ldxs[i];
    }
    if (dtype == MSX_BF16) return sum_launch_bf16<false>(a, 0, 1, 1, out, ldout, pixels, channels, nullptr, nullptr, nullptr, 0, BnFinalize{}, nullptr, as_stream(stream));
    MSX_DISPATCH_DTYPE(dtype, T, {
        const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
        { ProfScope ps__(K_SUM, as_stream(stream), (double)pixels * channels * (double)sizeof(T) * (n_in + 1.0), 0); sum_stats_kernel<T, false><<<dim3(stream_grid(pixels, c.lanes), c.nchunk), 256, 0, as_stream(stream)>>>(
            a, 0, 1, 1, static_cast<T*>(out), ldout, pixels, c.cp, c.vch, nullptr); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_sum_act(int dtype, int n_in, const void* const* host_xs, const int32_t* host_ldxs, int up2x_first,
                int32_t n, int32_t h, int32_t w, int32_t channels, int relu, void* out, int32_t ldout, void* stream) {
    MSX_REQUIRE(n_in >= 1 && n_in <= 8 && host_xs && host_ldxs && out, "sum_act: bad arguments");
    MSX_REQUIRE(!up2x_first || (h % 2 == 0 && w % 2 == 0), "sum_act: nearest-2x needs even h, w (got %dx%d)", h, w);
    MSX_TRY(check_ld(ldout, channels, "sum_act out"));
    SumArgs a;
    a.n = n_in;
    for (int i = 0; i < n_in; ++i) {
        MSX_TRY(check_ld(host_ldxs[i], channels, "sum_act x"));
        a.x[i] = host_xs[i];
        a.ld[i] = host_ldxs[i];
    }
    const int64_t pixels = (int64_t)n * h * w;
    if (dtype == MSX_BF16) return sum_launch_bf16<false>(a, up2x_first, h, w, out, ldout, pixels, channels, nullptr, nullptr, nullptr, relu, BnFinalize{}, nullptr, as_stream(stream));
    MSX_DISPATCH_DTYPE(dtype, T, {
        const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
        { ProfScope ps__(K_SUM, as_stream(stream), (double)pixels * channels * (double)sizeof(T) * (n_in + 1.0), 0); sum_stats_kernel<T, false><<<dim3(stream_grid(pixels, c.lanes), c.nchunk), 256, 0, as_stream(stream)>>>(
            a, up2x_first, h, w, static_cast<T*>(out), ldout, pixels, c.cp, c.vch, nullptr, nullptr, nullptr, relu); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_bn_act_bwd(int dtype, const void* s, int32_t lds, const void* dz, int32_t lddz, int64_t pixels, int32_t channels,
                   const float* gamma, const float* beta, const float* save_mean, const float* save_invstd, int relu,
                   void* ds, int32_t ldds, float* dgamma, float* dbeta, void* ws, size_t ws_bytes, void* stream) {
    MSX_REQUIRE(s && dz && ds && gamma && beta && save_mean && save_invstd, "bn_act_bwd: null pointer");
    return bn_bwd_launch(dtype, s, lds, dz, lddz, pixels, channels, gamma, beta, save_mean, save_invstd, relu, ds, ldds,
                         dgamma, dbeta, ws, ws_bytes, as_stream(stream));
}

// ---- SyncBatchNorm pieces (utils/parallel.py:37-38): the per-rank halves around the caller's all-reduce ----
int msx_bn_stats_sums(int dtype, const void* x, int32_t ldx, int64_t pixels, int32_t channels, double* sums, void* ws, size_t ws_bytes,
                      void* stream) {
    MSX_REQUIRE(x && sums && ws, "bn_stats_sums: null pointer");
    if (ws_bytes < bn_partials_bytes(channels)) return set_error(MSX_ERR_WORKSPACE, "bn_stats_sums: workspace %zu < %zu", ws_bytes, bn_partials_bytes(channels));
    BnFinalize f;
    MSX_TRY(bn_finalize_args(&f, static_cast<double*>(ws), channels, pixels, 0.f, 0.f, nullptr, nullptr, nullptr, nullptr, nullptr, sums));
    f.ticket = nullptr;
    return bn_stats_launch(dtype, x, ldx, pixels, channels, static_cast<double*>(ws), &f, nullptr, as_stream(stream));
}

int msx_bn_finalize_sums(const double* sums, int32_t channels, float eps, float momentum, float* save_mean, float* save_invstd,
                         float* running_mean, float* running_var, int64_t* num_batches_tracked, void* stream) {
    MSX_REQUIRE(sums && save_mean && save_invstd, "bn_finalize_sums: null pointer");
    BnFinalize f;
    MSX_TRY(bn_finalize_args(&f, nullptr, channels, 0, eps, momentum, save_mean, save_invstd, running_mean, running_var, num_batches_tracked, nullptr));
    f.ticket = nullptr;
    return bn_finalize_sums_launch(sums, f, as_stream(stream));
}

int msx_bn_apply(int dtype, const void* x, int32_t ldx, void* z, int32_t ldz, int64_t pixels, int32_t channels, const float* mean,
                 const float* invstd, const float* gamma, const float* beta, int relu, void* stream) {
    MSX_REQUIRE(x && z && mean && invstd && gamma && beta, "bn_apply: null pointer");
    return bn_apply_launch(dtype, x, ldx, z, ldz, pixels, channels, mean, invstd, gamma, beta, relu, as_stream(stream));
}

int msx_bn_bwd_reduce_sums(int dtype, const void* x, int32_t ldx, const void* dz, int32_t lddz, int64_t pixels, int32_t channels,
                           const float* gamma, const float* beta, const float* save_mean, const float* save_invstd, int relu,
                           float* dgamma, float* dbeta, double* sums, void* ws, size_t ws_bytes, void* stream) {
    MSX_REQUIRE(x && dz && gamma && beta && save_mean && save_invstd && sums && ws, "bn_bwd_reduce_sums: null pointer");
    if (ws_bytes < bn_partials_bytes(channels)) return set_error(MSX_ERR_WORKSPACE, "bn_bwd_reduce_sums: workspace %zu < %zu", ws_bytes, bn_partials_bytes(channels));
    const int cp = round_up(channels, 8);
    double* partials = static_cast<double*>(ws);
    float* c1 = reinterpret_cast<float*>(partials + (size_t)kMaxPartialBlocks * 2 * cp);
    return bn_bwd_reduce_launch(dtype, x, ldx, dz, lddz, pixels, channels, gamma, beta, save_mean, save_invstd, relu, dgamma, dbeta,
                                partials, c1, c1 + cp, sums, as_stream(stream));
}

int msx_bn_bwd_apply_sums(int dtype, const void* x, int32_t ldx, const void* dz, int32_t lddz, int64_t pixels, int32_t channels,
                          const float* gamma, const float* beta, const float* save_mean, const float* save_invstd, int relu,
                          const double* sums, const double* count_dev, void* dx, int32_t lddx, void* ws, size_t ws_bytes, void* stream) {
    MSX_REQUIRE(x && dz && dx && gamma && beta && save_mean && save_invstd && sums && count_dev && ws, "bn_bwd_apply_sums: null pointer");
    if (ws_bytes < bn_partials_bytes(channels)) return set_error(MSX_ERR_WORKSPACE, "bn_bwd_apply_sums: workspace %zu < %zu", ws_bytes, bn_partials_bytes(channels));
    const int cp = round_up(channels, 8);
    float* c1 = reinterpret_cast<float*>(static_cast<double*>(ws) + (size_t)kMaxPartialBlocks * 2 * cp);
    cudaStream_t st = as_stream(stream);
    { ProfScope ps__(K_BN_FINALIZE, st, 0, 0); bn_bwd_coeffs_kernel<<<(cp + 127) / 128, 128, 0, st>>>(sums, cp, channels, count_dev, c1, c1 + cp); }
    MSX_LAUNCH_CHECK();
    return bn_bwd_apply_launch(dtype, x, ldx, dz, lddz, pixels, channels, gamma, beta, save_mean, save_invstd, relu, c1, c1 + cp, dx, lddx, st);
}

int msx_upsample2x_bwd(int dtype, const void* ds, int32_t ldds, int32_t n, int32_t h, int32_t w, int32_t channels,
                       void* da, int32_t ldda, void* stream) {
    MSX_REQUIRE(ds && da, "upsample2x_bwd: null pointer");
    MSX_REQUIRE(h % 2 == 0 && w % 2 == 0, "upsample2x_bwd: odd size %dx%d", h, w);
    MSX_TRY(check_ld(ldds, channels, "upsample2x_bwd ds"));
    MSX_TRY(check_ld(ldda, channels, "upsample2x_bwd da"));
    const int64_t plo = (int64_t)n * (h / 2) * (w / 2);
    MSX_DISPATCH_DTYPE(dtype, T, {
        const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
        { ProfScope ps__(K_UPSAMPLE_BWD, as_stream(stream), (double)plo * channels * (double)sizeof(T) * 5.0, 0); upsample2x_bwd_kernel<T><<<dim3(stream_grid(plo, c.lanes), c.nchunk), 256, 0, as_stream(stream)>>>(
            static_cast<const T*>(ds), ldds, h, w, static_cast<T*>(da), ldda, plo, c.cp, c.vch); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_maxpool3s2_fwd(int dtype, const void* x, int32_t ldx, int32_t n, int32_t h, int32_t w, int32_t channels,
                       void* z, int32_t ldz, int8_t* idx, int32_t ldidx, void* stream) {
    MSX_REQUIRE(x && z && idx, "maxpool_fwd: null pointer");
    MSX_TRY(check_ld(ldx, channels, "maxpool x"));
    MSX_TRY(check_ld(ldz, channels, "maxpool z"));
    MSX_TRY(check_ld(ldidx, channels, "maxpool idx"));
    const int oh = (h - 1) / 2 + 1, ow = (w - 1) / 2 + 1;
    const int64_t opix = (int64_t)n * oh * ow;
    MSX_DISPATCH_DTYPE(dtype, T, {
        const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
        { ProfScope ps__(K_MAXPOOL, as_stream(stream), (double)opix * channels * ((double)sizeof(T) * 5.0 + 1.0), 0); maxpool_fwd_kernel<T><<<dim3(stream_grid(opix, c.lanes), c.nchunk), 256, 0, as_stream(stream)>>>(
            static_cast<const T*>(x), ldx, h, w, oh, ow, static_cast<T*>(z), ldz, idx, ldidx, opix, c.cp, c.vch); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_maxpool3s2_bwd(int dtype, const void* dz, int32_t lddz, const int8_t* idx, int32_t ldidx, int32_t n, int32_t h,
                       int32_t w, int32_t channels, void* dx, int32_t lddx, void* stream) {
    MSX_REQUIRE(dz && dx && idx, "maxpool_bwd: null pointer");
    MSX_TRY(check_ld(lddz, channels, "maxpool dz"));
    MSX_TRY(check_ld(lddx, channels, "maxpool dx"));
    const int oh = (h - 1) / 2 + 1, ow = (w - 1) / 2 + 1;
    const int64_t ipix = (int64_t)n * h * w;
    MSX_DISPATCH_DTYPE(dtype, T, {
        const ChanCfg c = chan_cfg<Vec<T>::N>(channels);
        { ProfScope ps__(K_MAXPOOL, as_stream(stream), (double)ipix * channels * ((double)sizeof(T) * 1.25 + 0.25 + (double)sizeof(T)), 0); maxpool_bwd_kernel<T><<<dim3(stream_grid(ipix, c.lanes), c.nchunk), 256, 0, as_stream(stream)>>>(
            static_cast<const T*>(dz), lddz, idx, ldidx, h, w, oh, ow, static_cast<T*>(dx), lddx, ipix, c.cp, c.vch); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_nchw_to_nhwc(int dtype, const float* src, int32_t n, int32_t c, int32_t h, int32_t w, void* dst, int32_t ld, void* stream) {
    MSX_REQUIRE(src && dst, "nchw_to_nhwc: null pointer");
    MSX_TRY(check_ld(ld, c, "nchw_to_nhwc"));
    const int64_t hw = (int64_t)h * w, pixels = hw * n;
    const int grid = (int)((pixels + 255) / 256 < kNumSMs * 8 ? (pixels + 255) / 256 : kNumSMs * 8);
    MSX_DISPATCH_DTYPE(dtype, T, {
        { ProfScope ps__(K_LAYOUT, as_stream(stream), (double)pixels * (c * 4.0 + round_up(c, 8) * (double)sizeof(T)), 0); nchw_to_nhwc_kernel<T><<<grid, 256, 0, as_stream(stream)>>>(src, c, hw, pixels, static_cast<T*>(dst), ld, round_up(c, 8)); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_nhwc_to_nchw(int dtype, const void* src, int32_t ld, int32_t n, int32_t c, int32_t h, int32_t w, float* dst, void* stream) {
    MSX_REQUIRE(src && dst, "nhwc_to_nchw: null pointer");
    const int64_t hw = (int64_t)h * w, pixels = hw * n;
    const int grid = (int)((pixels + 255) / 256 < kNumSMs * 8 ? (pixels + 255) / 256 : kNumSMs * 8);
    MSX_DISPATCH_DTYPE(dtype, T, {
        { ProfScope ps__(K_LAYOUT, as_stream(stream), (double)pixels * c * (4.0 + (double)sizeof(T)), 0); nhwc_to_nchw_kernel<T><<<grid, 256, 0, as_stream(stream)>>>(static_cast<const T*>(src), ld, c, hw, pixels, dst); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

}  // extern "C"
