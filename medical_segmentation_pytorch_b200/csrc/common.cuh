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

// ---- programmatic dependent launch ---------------------------------------------------------------------------------
// The step is a chain of ~2 800 short kernels; about a quarter of its time does not scale with the batch (launch hand-over,
// prologues, tails).  Kernels of the hot chain are launched with programmaticStreamSerialization: the next kernel of a stream
// may become resident while this one drains and runs its prologue (shared-memory carve-up, mbarrier init, TMEM allocation)
// up to pdl_wait(), which returns once every kernel before it has completed and its writes are visible.  A kernel launched
// this way must not touch global memory before pdl_wait() - neither read a predecessor's output nor write a buffer a
// predecessor may still read.  griddepcontrol.launch_dependents releases the successor's launch; a kernel that never issues it
// releases it at exit.
// MSX_PDL=0 launches the same kernels without the attribute (pdl_wait / pdl_trigger are then no-ops).
// Where the successor's launch is released (MSX_PDL_MODE, measured on the whole step, DUCKNet-17 batch 16):
//   2 (default)  never explicitly - at this kernel's exit: 409.4 img/s.  The gain is the pre-staged launch itself.
//   0            right after the wait: 404.0;   1  before the wait: 401.5 - successors that become resident early only sit
//                on registers and CTA slots that kernels of the sibling streams would have used.
//   3            pdl_trigger_late(): when a CTA has issued its last loads (producer warp) / finished its main loop
#ifndef MSX_PDL_MODE
#define MSX_PDL_MODE 2
#endif
__device__ __forceinline__ void pdl_wait() {
    if (MSX_PDL_MODE == 1) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}
__device__ __forceinline__ void pdl_trigger() {
    if (MSX_PDL_MODE == 0) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_trigger_late() {
    if (MSX_PDL_MODE == 3) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
int pdl_enabled();
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled();
    cfg.attrs = attr; cfg.numAttrs = 1;
    (void)cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);      // errors surface through MSX_LAUNCH_CHECK()
}

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


// ---- "the last CTA to arrive finalizes" ------------------------------------------------------------------------
// Channel reductions (BatchNorm statistics, BN-backward sums) are two-stage: every CTA writes one row of per-channel
// partials, then somebody folds the rows.  That second stage used to be its own launch (584 of them per DUCK-Net
// step, 9 us of pure latency each); now every CTA takes a ticket after its row is visible and the CTA that draws the
// last one does the fold.  Tickets live in a zero-initialised __device__ array of the library (api.cu), one slot per
// launch handed out round-robin by the host (a slot re-arms itself, so graph replays reuse theirs); two launches could
// only share a slot while both are in flight if kTicketSlots other reductions were launched in between.
// No allocation, no memset node, deterministic result (the fold order is fixed).
constexpr int kTicketSlots = 8192;
unsigned int* next_ticket();     // api.cu: device address of the next slot on the current device (nullptr on a CUDA error)

// True in every thread of the last of `total` CTAs to call it (all threads of the CTA must call).  `flag`: one int of
// shared memory.  The caller's global writes made before the call are visible to the last CTA after it.
__device__ inline bool ticket_is_last(unsigned int* ticket, unsigned int total, int* flag) {
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int t = atomicAdd(ticket, 1u);
        const int last = t == total - 1;
        if (last) *ticket = 0;                  // re-arm for the next launch that is handed this slot
        *flag = last;
    }
    __syncthreads();
    const bool last = *flag != 0;
    if (last) __threadfence();
    return last;
}

// Fold per-CTA partial rows [nblk][2][cp] (doubles) of channel c: one warp, fixed order.  L2 loads (__ldcg): the rows
// were written by other CTAs of the same launch.
__device__ inline void fold_partials(const double* partials, int nblk, int cp, int c, double& s, double& q) {
    const int lane = threadIdx.x & 31;
    s = 0.0; q = 0.0;
    int b = lane;
    for (; b + 96 < nblk; b += 128) {                           // 8 independent loads in flight
        double s4[4], q4[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            s4[u] = __ldcg(partials + ((int64_t)(b + 32 * u) * 2 + 0) * cp + c);
            q4[u] = __ldcg(partials + ((int64_t)(b + 32 * u) * 2 + 1) * cp + c);
        }
        s += (s4[0] + s4[1]) + (s4[2] + s4[3]);
        q += (q4[0] + q4[1]) + (q4[2] + q4[3]);
    }
    for (; b < nblk; b += 32) {
        s += __ldcg(partials + ((int64_t)b * 2 + 0) * cp + c);
        q += __ldcg(partials + ((int64_t)b * 2 + 1) * cp + c);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
    }
}

// What the forward statistics become (nn.BatchNorm2d, training): mean / invstd for this step, running statistics with
// the unbiased variance, num_batches_tracked.  With `sums` set nothing is finalized: the per-channel [sum, sum of
// squares] (doubles, [2][cp]) are left there for a cross-rank reduction (SyncBatchNorm) and msx_bn_finalize_sums.
struct BnFinalize {
    double* partials;            // [rows][2][cp]
    unsigned int* ticket;        // nullptr: no in-kernel finalize
    double* sums;                // nullable
    float *mean, *invstd, *running_mean, *running_var;
    long long* nbt;
    long long count;             // pixels per channel
    int channels, cp;
    float eps, momentum;
};

__device__ inline void bn_finalize_one(const BnFinalize& f, int c, double s, double q) {
    const double m = s / (double)f.count;
    double var = q / (double)f.count - m * m;
    if (var < 0.0) var = 0.0;
    f.mean[c] = (float)m;
    f.invstd[c] = (float)(1.0 / sqrt(var + (double)f.eps));
    if (f.running_mean) {
        const double unb = f.count > 1 ? var * ((double)f.count / (double)(f.count - 1)) : var;
        f.running_mean[c] = (float)((1.0 - f.momentum) * (double)f.running_mean[c] + (double)f.momentum * m);
        f.running_var[c] = (float)((1.0 - f.momentum) * (double)f.running_var[c] + (double)f.momentum * unb);
    }
}

// Executed by every thread of the last CTA: warp w handles channels w, w + nwarps, ...
__device__ inline void bn_finalize_all(const BnFinalize& f, int nblk) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
    if (threadIdx.x == 0 && f.nbt && !f.sums) *f.nbt += 1;
    for (int c = warp; c < f.cp; c += nwarps) {
        if (c >= f.channels) {
            if (lane == 0) {
                if (f.sums) { f.sums[c] = 0.0; f.sums[f.cp + c] = 0.0; }
                else { f.mean[c] = 0.f; f.invstd[c] = 0.f; }
            }
            continue;
        }
        double s, q;
        fold_partials(f.partials, nblk, f.cp, c, s, q);
        if (lane != 0) continue;
        if (f.sums) { f.sums[c] = s; f.sums[f.cp + c] = q; }
        else bn_finalize_one(f, c, s, q);
    }
}

// What a gather-form conv kernel does to an accumulator column c before storing it:
//   v += bias[c] (ConvTranspose2d bias);  v = v * scale[c] + shift[c] (eval-mode BatchNorm folded to an affine);  ReLU.
// All pointers nullable; columns >= the logical channel count store zeros.
//   With `stats` the kernel also produces the training-mode BatchNorm statistics of what it stores (sum and sum of
//   squares per channel of the bf16-rounded outputs) and its last CTA finalizes them: no separate pass over y.
struct ConvEpilogue {
    const float* bias;
    const float* scale;
    const float* shift;
    int relu;
    int stats;                   // 0 / 1: bnf is valid
    BnFinalize bnf;
};
inline ConvEpilogue conv_epilogue(const float* bias, const float* scale = nullptr, const float* shift = nullptr, int relu = 0) {
    ConvEpilogue e;
    e.bias = bias; e.scale = scale; e.shift = shift; e.relu = relu; e.stats = 0;
    e.bnf = BnFinalize{};
    return e;
}

// ---- internal cross-file entry points -------------------------------------------------------
// channel-reduction partial sums: [nblk][2][cp] doubles
size_t bn_partials_bytes(int channels);
int bn_finalize_args(BnFinalize* f, double* partials, int channels, int64_t count, float eps, float momentum, float* mean,
                     float* invstd, float* running_mean, float* running_var, int64_t* nbt, double* sums);
int bn_stats_launch(int dtype, const void* x, int ld, int64_t pixels, int channels, double* partials, const BnFinalize* fin,
                    int* nblk_out, cudaStream_t st);
int bn_finalize_sums_launch(const double* sums, const BnFinalize& f, cudaStream_t st);
int bn_finalize_launch(const BnFinalize& f, int nblk, cudaStream_t st);
int ticket_max_cp();
int bn_bwd_reduce_launch(int dtype, const void* x, int ldx, const void* dz, int lddz, int64_t pixels, int channels,
                         const float* gamma, const float* beta, const float* mean, const float* invstd, int flags,
                         float* dgamma, float* dbeta, double* partials, float* c1, float* c2, double* sums, cudaStream_t st);
int bn_bwd_apply_launch(int dtype, const void* x, int ldx, const void* dz, int lddz, int64_t pixels, int channels,
                        const float* gamma, const float* beta, const float* mean, const float* invstd, int flags,
                        const float* c1, const float* c2, void* dx, int lddx, cudaStream_t st);
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
