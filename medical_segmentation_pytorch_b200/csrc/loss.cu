// Loss and metric kernels (tensor-core-free, HBM-bound).  Logits are NCHW fp32 exactly as the
// reference hands them to nn.CrossEntropyLoss / F.kl_div / torchmetrics; labels are int64.
// A thread owns PV consecutive pixels (PV = 4 when h*w % 4 == 0: one 16 B load per class plane and
// two 16 B label loads), keeps the <= CT class values in registers, and emits loss and gradient in
// the same pass.  Sums are accumulated per thread in fp32 over few elements, then in double.
#include "common.cuh"

#include <map>
#include <mutex>

namespace msx {
namespace {

constexpr int kLossMaxBlocks = kNumSMs * 8;   // 1184
// workspace layout (doubles): [0,1184) weight partials, [1184,2368) loss partials, then scalars
constexpr int kWsWeight = 0, kWsLoss = kLossMaxBlocks, kWsScalars = 2 * kLossMaxBlocks, kWsDoubles = 2 * kLossMaxBlocks + 64;

template <int PV>
__device__ inline void load_px(const float* p, float (&v)[PV]) {
    if constexpr (PV == 8) {
        float4 r = *reinterpret_cast<const float4*>(p), u = *reinterpret_cast<const float4*>(p + 4);
        v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w; v[4] = u.x; v[5] = u.y; v[6] = u.z; v[7] = u.w;
    } else if constexpr (PV == 4) { float4 r = *reinterpret_cast<const float4*>(p); v[0] = r.x; v[1] = r.y; v[2] = r.z; v[3] = r.w; }
    else v[0] = *p;
}
template <int PV>
__device__ inline void store_px(float* p, const float (&v)[PV]) {
    if constexpr (PV == 8) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
    } else if constexpr (PV == 4) *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    else *p = v[0];
}
template <int PV>
__device__ inline void load_lab(const int64_t* p, int64_t (&t)[PV]) {
    if constexpr (PV == 8) {
        longlong2 a = *reinterpret_cast<const longlong2*>(p), b = *reinterpret_cast<const longlong2*>(p + 2);
        longlong2 c = *reinterpret_cast<const longlong2*>(p + 4), d = *reinterpret_cast<const longlong2*>(p + 6);
        t[0] = a.x; t[1] = a.y; t[2] = b.x; t[3] = b.y; t[4] = c.x; t[5] = c.y; t[6] = d.x; t[7] = d.y;
    } else if constexpr (PV == 4) {
        longlong2 a = *reinterpret_cast<const longlong2*>(p), b = *reinterpret_cast<const longlong2*>(p + 2);
        t[0] = a.x; t[1] = a.y; t[2] = b.x; t[3] = b.y;
    } else t[0] = *p;
}

// One-byte label codes written by the weight pre-pass: the class index, or 255 for ignored / out-of-range
// pixels.  The main pass then reads 1 B/pixel instead of the 8 B int64 label a second time.
// Both loaders hand the kernel 32-bit class indices with -1 for "no contribution", so the per-class
// compares in the hot loop are single ISETPs (the int64 form made the kernel issue-bound).
template <int PV>
__device__ inline void load_codes(const uint8_t* p, int (&t)[PV]) {
    if constexpr (PV == 8) {
        const uint2 r = *reinterpret_cast<const uint2*>(p);
#pragma unroll
        for (int j = 0; j < 4; ++j) { t[j] = (r.x >> (8 * j)) & 255u; t[4 + j] = (r.y >> (8 * j)) & 255u; }
    } else if constexpr (PV == 4) {
        const uint32_t r = *reinterpret_cast<const uint32_t*>(p);
#pragma unroll
        for (int j = 0; j < 4; ++j) t[j] = (r >> (8 * j)) & 255u;
    } else t[0] = *p;
#pragma unroll
    for (int j = 0; j < PV; ++j) if (t[j] == 255) t[j] = -1;
}
template <int PV>
__device__ inline void load_lab_idx(const int64_t* p, int64_t ignore, int c, int (&ti)[PV]) {
    int64_t t[PV];
    load_lab<PV>(p, t);
#pragma unroll
    for (int j = 0; j < PV; ++j) ti[j] = (t[j] != ignore && (uint64_t)t[j] < (uint64_t)c) ? (int)t[j] : -1;
}
// MUFU-based exp / log / reciprocal.  Arguments of the exp are <= 0 (max-subtracted), so the absolute error
// of a softmax probability is < 3e-8; lg2.approx is within 2^-22 absolute on the sum (>= 1).
// The .ftz forms skip the denormal fix-up code of __expf/__logf: results below 2^-126 flush to zero, the
// arguments of log and rcp are sums >= 1.
__device__ inline float fast_exp(float x) {
    float r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x * 1.4426950408889634f));
    return r;
}
__device__ inline float fast_log(float x) {
    float r;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r * 0.6931471805599453f;
}
__device__ inline float fast_rcp(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
template <int PV>
__device__ inline void store_codes(uint8_t* p, const uint32_t (&cd)[PV]) {
    if constexpr (PV == 4) *reinterpret_cast<uint32_t*>(p) = cd[0] | (cd[1] << 8) | (cd[2] << 16) | (cd[3] << 24);
    else p[0] = (uint8_t)cd[0];
}

__device__ inline double block_sum_d(double v) {
    __shared__ double sh[32];
    return block_sum<double>(v, sh);
}

// ---- cross entropy --------------------------------------------------------------------------
template <int PV>
__global__ void __launch_bounds__(256) ce_wsum_kernel(const int64_t* __restrict__ labels, int64_t groups, int c,
                                                      const float* __restrict__ cw, int64_t ignore, double* __restrict__ ws,
                                                      uint8_t* __restrict__ codes) {
    float acc = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < groups; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t t[PV];
        uint32_t cd[PV];
        load_lab<PV>(labels + i * PV, t);
#pragma unroll
        for (int j = 0; j < PV; ++j) {
            const bool valid = t[j] != ignore && t[j] >= 0 && t[j] < c;
            if (valid) acc += cw ? cw[t[j]] : 1.f;
            cd[j] = valid ? (uint32_t)t[j] : 255u;
        }
        if (codes) store_codes<PV>(codes + i * PV, cd);
    }
    const double s = block_sum_d((double)acc);
    if (threadIdx.x == 0) ws[kWsWeight + blockIdx.x] = s;
}

// MODE 0: fused mean/sum loss + gradient.  MODE 1: per-pixel loss only ('none' forward).
// MODE 2: gradient with a per-pixel upstream coefficient ('none' backward / OHEM).
template <int PV, int CT, int MODE>
__global__ void __launch_bounds__(256) ce_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels,
                                                 int c, int64_t hw, int64_t groups, const float* __restrict__ cw, int64_t ignore,
                                                 int nwpart, int use_denom, const float* __restrict__ coef_pix,
                                                 float* __restrict__ out_pix, float* __restrict__ dlogits, double* __restrict__ ws,
                                                 const uint8_t* __restrict__ codes) {
    double inv_denom = 1.0;
    if (MODE == 0 && use_denom) {
        // every CTA re-reduces the (<= 1184) weight partials: cheaper than one more launch
        double d = 0.0;
        for (int i = threadIdx.x; i < nwpart; i += blockDim.x) d += ws[kWsWeight + i];
        __shared__ double shd[32];
        __shared__ double tot;
        d = block_sum<double>(d, shd);
        if (threadIdx.x == 0) tot = d;
        __syncthreads();
        inv_denom = 1.0 / tot;
    }
    const float invd = (float)inv_denom;
    const int64_t gpi = hw / PV;   // pixel groups per image
    float lacc = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < groups; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t iu = (uint32_t)i, nu = iu / (uint32_t)gpi;      // 32-bit: groups < 2^31 (checked on the host)
        const int64_t n = nu, q = (int64_t)(iu - nu * (uint32_t)gpi) * PV;
        const float* zp = logits + n * c * hw + q;
        int t[PV];
        if (codes) load_codes<PV>(codes + n * hw + q, t);
        else load_lab_idx<PV>(labels + n * hw + q, ignore, c, t);
        float z[CT > 0 ? CT : 1][PV];
        float mx[PV], se[PV], zt[PV];
#pragma unroll
        for (int j = 0; j < PV; ++j) { mx[j] = -INFINITY; se[j] = 0.f; zt[j] = 0.f; }
        if constexpr (CT > 0) {
#pragma unroll
            for (int k = 0; k < CT; ++k)
                if (k < c) {
                    load_px<PV>(zp + k * hw, z[k]);
#pragma unroll
                    for (int j = 0; j < PV; ++j) mx[j] = fmaxf(mx[j], z[k][j]);
                }
#pragma unroll
            for (int k = 0; k < CT; ++k)
                if (k < c) {
#pragma unroll
                    for (int j = 0; j < PV; ++j) {
                        if (t[j] == k) zt[j] = z[k][j];
                        z[k][j] = fast_exp(z[k][j] - mx[j]);      // z now holds exp(z - max): reused by the gradient
                        se[j] += z[k][j];
                    }
                }
        } else {
            for (int k = 0; k < c; ++k) {
                float v[PV];
                load_px<PV>(zp + k * hw, v);
#pragma unroll
                for (int j = 0; j < PV; ++j) mx[j] = fmaxf(mx[j], v[j]);
            }
            for (int k = 0; k < c; ++k) {
                float v[PV];
                load_px<PV>(zp + k * hw, v);
#pragma unroll
                for (int j = 0; j < PV; ++j) {
                    se[j] += fast_exp(v[j] - mx[j]);
                    if (t[j] == k) zt[j] = v[j];
                }
            }
        }
        float wt[PV], lp[PV], up[PV];
        if (MODE == 2) load_px<PV>(coef_pix + n * hw + q, up);
#pragma unroll
        for (int j = 0; j < PV; ++j) {
            wt[j] = t[j] >= 0 ? (cw ? cw[t[j]] : 1.f) : 0.f;
            lp[j] = t[j] >= 0 ? wt[j] * ((mx[j] + fast_log(se[j])) - zt[j]) : 0.f;
            lacc += lp[j];
            if (MODE == 0) up[j] = invd;
        }
        if (MODE == 1) { store_px<PV>(out_pix + n * hw + q, lp); continue; }
        float rse[PV];
#pragma unroll
        for (int j = 0; j < PV; ++j) { rse[j] = fast_rcp(se[j]); wt[j] = t[j] >= 0 ? wt[j] * up[j] : 0.f; }   // exactly 0 on ignored pixels
        if (dlogits) {
            float* gp = dlogits + n * c * hw + q;
            if constexpr (CT > 0) {
#pragma unroll
                for (int k = 0; k < CT; ++k)
                    if (k < c) {
                        float g[PV];
#pragma unroll
                        for (int j = 0; j < PV; ++j) {
                            const float sm = z[k][j] * rse[j];
                            g[j] = wt[j] * (t[j] == k ? sm - 1.f : sm);
                        }
                        store_px<PV>(gp + k * hw, g);
                    }
            } else {
                for (int k = 0; k < c; ++k) {
                    float v[PV], g[PV];
                    load_px<PV>(zp + k * hw, v);
#pragma unroll
                    for (int j = 0; j < PV; ++j) {
                        const float sm = fast_exp(v[j] - mx[j]) * rse[j];
                        g[j] = wt[j] * (t[j] == k ? sm - 1.f : sm);
                    }
                    store_px<PV>(gp + k * hw, g);
                }
            }
        }
    }
    if (MODE == 0) {
        const double s = block_sum_d((double)lacc);
        if (threadIdx.x == 0) ws[kWsLoss + blockIdx.x] = s;
    }
}

__global__ void ce_finalize_kernel(const double* __restrict__ ws, int nwpart, int nlpart, int use_denom, float* loss_out) {
    __shared__ double sh[32];
    double l = 0.0, w = 0.0;
    for (int i = threadIdx.x; i < nlpart; i += blockDim.x) l += ws[kWsLoss + i];
    for (int i = threadIdx.x; i < nwpart; i += blockDim.x) w += ws[kWsWeight + i];
    l = block_sum<double>(l, sh);
    w = block_sum<double>(w, sh);
    if (threadIdx.x == 0) *loss_out = use_denom ? (float)(l / w) : (float)l;
}

inline bool groups_ok(int64_t groups) { return groups < (int64_t)1 << 31; }
inline int loss_grid(int64_t groups) {
    int64_t g = cdiv64(groups, 256);
    if (g < 1) g = 1;
    if (g > kLossMaxBlocks) g = kLossMaxBlocks;
    return (int)g;
}

// Grid = exactly the CTAs that are resident at once (SMs x occupancy), each striding over the pixel groups:
// with a fixed 8 CTAs/SM the register-heavy variants (3 resident) ran 2.67 waves and lost the last third.
int resident_grid(const void* kernel, size_t smem, int64_t groups) {
    static std::mutex mu;
    static std::map<std::pair<const void*, size_t>, int> cache;
    int occ = 0;
    {
        std::lock_guard<std::mutex> lk(mu);
        auto it = cache.find({kernel, smem});
        if (it != cache.end()) occ = it->second;
        else {
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, 256, smem) != cudaSuccess || occ < 1) { (void)cudaGetLastError(); occ = 1; }
            if (occ > 8) occ = 8;
            cache[{kernel, smem}] = occ;
        }
    }
    int64_t g = cdiv64(groups, 256);
    if (g < 1) g = 1;
    if (g > (int64_t)kNumSMs * occ) g = (int64_t)kNumSMs * occ;
    return (int)g;
}

template <int MODE>
int ce_dispatch(const float* logits, const int64_t* labels, int n, int c, int64_t hw, const float* cw, int64_t ignore,
                int nwpart, int use_denom, const float* coef, float* out_pix, float* dlogits, double* ws, int* grid_out,
                cudaStream_t st, const uint8_t* codes = nullptr) {
    const bool v8 = hw % 8 == 0 && c <= 4, v4 = hw % 4 == 0;
    const int64_t groups = (int64_t)n * hw / (v8 ? 8 : (v4 ? 4 : 1));
    if (!groups_ok(groups)) return set_error(MSX_ERR_UNSUPPORTED, "loss: more than 2^31 pixel groups");
    int g = 1;
#define MSX_CE_LAUNCH(PV, CT) { g = resident_grid((const void*)ce_kernel<PV, CT, MODE>, 0, groups); ProfScope ps__(K_LOSS, st, (double)n * hw * (MODE == 1 ? 4.0 * c + 12.0 : (MODE == 2 ? 8.0 * c + 12.0 : (dlogits ? 8.0 * c + 8.0 : 4.0 * c + 8.0))), 0); ce_kernel<PV, CT, MODE><<<g, 256, 0, st>>>(logits, labels, c, hw, groups, cw, ignore, nwpart, use_denom, coef, out_pix, dlogits, ws, codes); }
    if (v8) {
        if (c <= 2) MSX_CE_LAUNCH(8, 2) else MSX_CE_LAUNCH(8, 4)
    } else if (v4) {
        if (c <= 2) MSX_CE_LAUNCH(4, 2) else if (c <= 4) MSX_CE_LAUNCH(4, 4) else if (c <= 8) MSX_CE_LAUNCH(4, 8) else MSX_CE_LAUNCH(4, 0)
    } else {
        if (c <= 2) MSX_CE_LAUNCH(1, 2) else if (c <= 4) MSX_CE_LAUNCH(1, 4) else if (c <= 8) MSX_CE_LAUNCH(1, 8) else MSX_CE_LAUNCH(1, 0)
    }
#undef MSX_CE_LAUNCH
    *grid_out = g;
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

// ---- knowledge distillation -----------------------------------------------------------------
// Register-resident variant for c <= CT classes: every logit is loaded once, 2c exp + 2 log per pixel.
template <int PV, int CT>
__global__ void __launch_bounds__(256) kd_kl_reg_kernel(const float* __restrict__ s, const float* __restrict__ t, int c, int64_t hw,
                                                        int64_t groups, float invT, float gscale, float* __restrict__ ds,
                                                        double* __restrict__ ws) {
    const int64_t gpi = hw / PV;
    float lacc = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < groups; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t iu = (uint32_t)i, nu = iu / (uint32_t)gpi;      // 32-bit: groups < 2^31 (checked on the host)
        const int64_t n = nu, q = (int64_t)(iu - nu * (uint32_t)gpi) * PV;
        const float* sp = s + n * c * hw + q;
        const float* tp = t + n * c * hw + q;
        float a[CT][PV], b[CT][PV], ms[PV], mt[PV], es[PV], et[PV];
#pragma unroll
        for (int j = 0; j < PV; ++j) { ms[j] = mt[j] = -INFINITY; es[j] = et[j] = 0.f; }
#pragma unroll
        for (int k = 0; k < CT; ++k)
            if (k < c) {
                load_px<PV>(sp + k * hw, a[k]);
                load_px<PV>(tp + k * hw, b[k]);
            }
#pragma unroll
        for (int k = 0; k < CT; ++k)
            if (k < c) {
#pragma unroll
                for (int j = 0; j < PV; ++j) {
                    a[k][j] *= invT; b[k][j] *= invT;
                    ms[j] = fmaxf(ms[j], a[k][j]); mt[j] = fmaxf(mt[j], b[k][j]);
                }
            }
        // a, b become the max-shifted logits; ea, eb their exponentials (kept for softmax and gradient)
        float ea[CT][PV], eb[CT][PV];
#pragma unroll
        for (int k = 0; k < CT; ++k)
            if (k < c) {
#pragma unroll
                for (int j = 0; j < PV; ++j) {
                    a[k][j] -= ms[j]; b[k][j] -= mt[j];
                    ea[k][j] = fast_exp(a[k][j]); eb[k][j] = fast_exp(b[k][j]);
                    es[j] += ea[k][j]; et[j] += eb[k][j];
                }
            }
        float dl[PV], rs[PV], rt[PV];      // dl = log(sum_s) - log(sum_t): logq - logp = (b - a) + dl
#pragma unroll
        for (int j = 0; j < PV; ++j) { dl[j] = fast_log(es[j]) - fast_log(et[j]); rs[j] = fast_rcp(es[j]) * gscale; rt[j] = fast_rcp(et[j]); }
#pragma unroll
        for (int k = 0; k < CT; ++k)
            if (k < c) {
                float g[PV];
#pragma unroll
                for (int j = 0; j < PV; ++j) {
                    const float qq = eb[k][j] * rt[j];
                    if (qq > 0.f) lacc = fmaf(qq, (b[k][j] - a[k][j]) + dl[j], lacc);
                    g[j] = fmaf(ea[k][j], rs[j], -qq * gscale);
                }
                if (ds) store_px<PV>(ds + n * c * hw + q + k * hw, g);
            }
    }
    const double r = block_sum_d((double)lacc);
    if (threadIdx.x == 0) ws[kWsLoss + blockIdx.x] = r;
}

template <int PV>
__global__ void __launch_bounds__(256) kd_kl_kernel(const float* __restrict__ s, const float* __restrict__ t, int c, int64_t hw,
                                                    int64_t groups, float invT, float gscale, float* __restrict__ ds,
                                                    double* __restrict__ ws) {
    const int64_t gpi = hw / PV;
    float lacc = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < groups; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t iu = (uint32_t)i, nu = iu / (uint32_t)gpi;      // 32-bit: groups < 2^31 (checked on the host)
        const int64_t n = nu, q = (int64_t)(iu - nu * (uint32_t)gpi) * PV;
        const float* sp = s + n * c * hw + q;
        const float* tp = t + n * c * hw + q;
        float ms[PV], mt[PV], es[PV], et[PV];
#pragma unroll
        for (int j = 0; j < PV; ++j) { ms[j] = mt[j] = -INFINITY; es[j] = et[j] = 0.f; }
        for (int k = 0; k < c; ++k) {
            float a[PV], b[PV];
            load_px<PV>(sp + k * hw, a);
            load_px<PV>(tp + k * hw, b);
#pragma unroll
            for (int j = 0; j < PV; ++j) { ms[j] = fmaxf(ms[j], a[j] * invT); mt[j] = fmaxf(mt[j], b[j] * invT); }
        }
        for (int k = 0; k < c; ++k) {
            float a[PV], b[PV];
            load_px<PV>(sp + k * hw, a);
            load_px<PV>(tp + k * hw, b);
#pragma unroll
            for (int j = 0; j < PV; ++j) { es[j] += expf(a[j] * invT - ms[j]); et[j] += expf(b[j] * invT - mt[j]); }
        }
        for (int k = 0; k < c; ++k) {
            float a[PV], b[PV], g[PV];
            load_px<PV>(sp + k * hw, a);
            load_px<PV>(tp + k * hw, b);
#pragma unroll
            for (int j = 0; j < PV; ++j) {
                const float logp = a[j] * invT - ms[j] - logf(es[j]);
                const float logq = b[j] * invT - mt[j] - logf(et[j]);
                const float qq = expf(logq);
                if (qq > 0.f) lacc += qq * (logq - logp);
                g[j] = (expf(logp) - qq) * gscale;
            }
            if (ds) store_px<PV>(ds + n * c * hw + q + k * hw, g);
        }
    }
    const double r = block_sum_d((double)lacc);
    if (threadIdx.x == 0) ws[kWsLoss + blockIdx.x] = r;
}

__global__ void __launch_bounds__(256) kd_mse_kernel(const float* __restrict__ s, const float* __restrict__ t, int64_t count,
                                                     float gscale, float* __restrict__ ds, double* __restrict__ ws) {
    float lacc = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const float d = s[i] - t[i];
        lacc += d * d;
        if (ds) ds[i] = d * gscale;
    }
    const double r = block_sum_d((double)lacc);
    if (threadIdx.x == 0) ws[kWsLoss + blockIdx.x] = r;
}

__global__ void scaled_sum_finalize_kernel(const double* __restrict__ ws, int npart, double scale, float* out) {
    __shared__ double sh[32];
    double l = 0.0;
    for (int i = threadIdx.x; i < npart; i += blockDim.x) l += ws[kWsLoss + i];
    l = block_sum<double>(l, sh);
    if (threadIdx.x == 0) *out = (float)(l * scale);
}

// ---- sigmoid + dice + bce ---------------------------------------------------------------------
template <bool FLOAT_T>
__device__ inline float target_at(const void* t, int64_t i) {
    return FLOAT_T ? static_cast<const float*>(t)[i] : (float)static_cast<const int64_t*>(t)[i];
}
template <bool FLOAT_T>
__device__ inline void targets4(const void* t, int64_t i, float (&y)[4]) {
    if (FLOAT_T) load_px<4>(static_cast<const float*>(t) + i, y);
    else {
        int64_t l[4];
        load_lab<4>(static_cast<const int64_t*>(t) + i, l);
#pragma unroll
        for (int j = 0; j < 4; ++j) y[j] = (float)l[j];
    }
}
// partial sums: 4 arrays of kLossMaxBlocks doubles starting at ws[0]
template <bool FLOAT_T>
__global__ void __launch_bounds__(256) dice_bce_reduce_kernel(const float* __restrict__ z, const void* __restrict__ t, int64_t count,
                                                              double* __restrict__ ws) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    auto body = [&](float x, float y) {
        const float e = fast_exp(-fabsf(x));                    // one exp serves softplus and sigmoid
        const float sp = fmaxf(x, 0.f) + fast_log(1.f + e);     // 1 + e in [1, 2]: absolute error < 2e-7
        const float r = fast_rcp(1.f + e);
        const float p = x >= 0.f ? r : e * r;
        a0 += sp - x * y; a1 += p * y; a2 += p; a3 += y;
    };
    const int64_t quads = count >> 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < quads; i += (int64_t)gridDim.x * blockDim.x) {
        float x[4], y[4];
        load_px<4>(z + 4 * i, x);
        targets4<FLOAT_T>(t, 4 * i, y);
#pragma unroll
        for (int j = 0; j < 4; ++j) body(x[j], y[j]);
    }
    for (int64_t i = 4 * quads + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
        body(z[i], target_at<FLOAT_T>(t, i));
    __shared__ double sh[32];
    const double r0 = block_sum<double>((double)a0, sh), r1 = block_sum<double>((double)a1, sh);
    const double r2 = block_sum<double>((double)a2, sh), r3 = block_sum<double>((double)a3, sh);
    if (threadIdx.x == 0) {
        ws[0 * kLossMaxBlocks / 2 + blockIdx.x] = r0;
        ws[1 * kLossMaxBlocks / 2 + blockIdx.x] = r1;
        ws[2 * kLossMaxBlocks / 2 + blockIdx.x] = r2;
        ws[3 * kLossMaxBlocks / 2 + blockIdx.x] = r3;
    }
}
__global__ void dice_bce_finalize_kernel(double* __restrict__ ws, int npart, int64_t count, float eps, float* loss_out) {
    __shared__ double sh[32];
    double r[4];
    for (int k = 0; k < 4; ++k) {
        double l = 0.0;
        for (int i = threadIdx.x; i < npart; i += blockDim.x) l += ws[k * kLossMaxBlocks / 2 + i];
        r[k] = block_sum<double>(l, sh);
    }
    if (threadIdx.x == 0) {
        const double num = 2.0 * r[1] + eps, den = r[2] + r[3] + eps;
        *loss_out = (float)(r[0] / (double)count + 1.0 - num / den);
        ws[kWsScalars + 0] = num;
        ws[kWsScalars + 1] = den;
    }
}
template <bool FLOAT_T>
__global__ void __launch_bounds__(256) dice_bce_grad_kernel(const float* __restrict__ z, const void* __restrict__ t, int64_t count,
                                                            const double* __restrict__ ws, float* __restrict__ dz) {
    const float num = (float)ws[kWsScalars + 0], den = (float)ws[kWsScalars + 1];
    const float inv_cnt = 1.f / (float)count, inv_den2 = 1.f / (den * den);
    auto grad = [&](float x, float y) {
        const float e = fast_exp(-fabsf(x)), r = fast_rcp(1.f + e);
        const float p = x >= 0.f ? r : e * r;
        const float ddice_dp = -(2.f * y * den - num) * inv_den2;
        return (p - y) * inv_cnt + ddice_dp * p * (1.f - p);
    };
    const int64_t quads = count >> 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < quads; i += (int64_t)gridDim.x * blockDim.x) {
        float x[4], y[4], g[4];
        load_px<4>(z + 4 * i, x);
        targets4<FLOAT_T>(t, 4 * i, y);
#pragma unroll
        for (int j = 0; j < 4; ++j) g[j] = grad(x[j], y[j]);
        store_px<4>(dz + 4 * i, g);
    }
    for (int64_t i = 4 * quads + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x)
        dz[i] = grad(z[i], target_at<FLOAT_T>(t, i));
}

__global__ void scale_kernel(float* __restrict__ d, int64_t count, const float* __restrict__ sc) {
    const float s = *sc;
    if (s == 1.f) return;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) d[i] *= s;
}

// ---- confusion matrix -----------------------------------------------------------------------
template <int PV>
__global__ void __launch_bounds__(256) confmat_kernel(const float* __restrict__ logits, const int64_t* __restrict__ labels, int c,
                                                      int64_t hw, int64_t groups, int64_t ignore,
                                                      unsigned long long* __restrict__ confmat, int* __restrict__ bad) {
    extern __shared__ unsigned int cnt[];   // c<=4: [c*c][256] private counters; else [c*c] shared atomics
    const bool priv = c <= 4;
    const int cells = c * c;
    if (priv) { for (int k = 0; k < cells; ++k) cnt[k * 256 + threadIdx.x] = 0; }
    else { for (int k = threadIdx.x; k < cells; k += blockDim.x) cnt[k] = 0; }
    __syncthreads();
    const int64_t gpi = hw / PV;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < groups; i += (int64_t)gridDim.x * blockDim.x) {
        const uint32_t iu = (uint32_t)i, nu = iu / (uint32_t)gpi;      // 32-bit: groups < 2^31 (checked on the host)
        const int64_t n = nu, q = (int64_t)(iu - nu * (uint32_t)gpi) * PV;
        const float* zp = logits + n * c * hw + q;
        int64_t t[PV];
        load_lab<PV>(labels + n * hw + q, t);
        float best[PV];
        int bi[PV];
        load_px<PV>(zp, best);
#pragma unroll
        for (int j = 0; j < PV; ++j) bi[j] = 0;
        for (int k = 1; k < c; ++k) {
            float v[PV];
            load_px<PV>(zp + k * hw, v);
#pragma unroll
            for (int j = 0; j < PV; ++j) {
                // torch.argmax: first maximal index; NaN is maximal (first NaN wins)
                const bool take = (v[j] > best[j]) || (v[j] != v[j] && best[j] == best[j]);
                if (take) { best[j] = v[j]; bi[j] = k; }
            }
        }
#pragma unroll
        for (int j = 0; j < PV; ++j) {
            if (t[j] == ignore) continue;
            if (t[j] < 0 || t[j] >= c) { *bad = 1; continue; }
            const int cell = (int)t[j] * c + bi[j];
            if (priv) cnt[cell * 256 + threadIdx.x] += 1; else atomicAdd(&cnt[cell], 1u);
        }
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

// ---- OHEM (device-side branch + radix select) -----------------------------------------------
// scalars in ws (doubles at kWsScalars..): see enum
enum { kOhValid = 0, kOhHard = 1, kOhHardSum = 2, kOhMode = 3, kOhK = 4, kOhPrefix = 5, kOhRemain = 6, kOhV = 7, kOhCntGt = 8, kOhCntEq = 9, kOhSumGt = 10 };

__device__ inline unsigned int f2ord(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ inline float ord2f(unsigned int o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

__global__ void __launch_bounds__(256) ohem_count_kernel(const float* __restrict__ loss_pix, const int64_t* __restrict__ labels,
                                                         int64_t count, int64_t ignore, float th, double* __restrict__ ws) {
    // partials: [0..) valid count, [1184/2..) hard count, loss partial region: hard sum
    float nv = 0.f, nh = 0.f, sh_ = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        if (labels[i] != ignore) nv += 1.f;
        const float l = loss_pix[i];
        if (l > th) { nh += 1.f; sh_ += l; }
    }
    __shared__ double sh[32];
    const double a = block_sum<double>((double)nv, sh), b = block_sum<double>((double)nh, sh), c = block_sum<double>((double)sh_, sh);
    if (threadIdx.x == 0) {
        ws[blockIdx.x] = a;
        ws[kLossMaxBlocks / 2 + blockIdx.x] = b;
        ws[kWsLoss + blockIdx.x] = c;
    }
}
__global__ void ohem_decide_kernel(double* __restrict__ ws, int npart, unsigned int* __restrict__ hist) {
    __shared__ double sh[32];
    double a = 0, b = 0, c = 0;
    for (int i = threadIdx.x; i < npart; i += blockDim.x) { a += ws[i]; b += ws[kLossMaxBlocks / 2 + i]; c += ws[kWsLoss + i]; }
    a = block_sum<double>(a, sh); b = block_sum<double>(b, sh); c = block_sum<double>(c, sh);
    for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
    if (threadIdx.x == 0) {
        double* S = ws + kWsScalars;
        const long long n_min = (long long)a / 16;
        S[kOhValid] = a; S[kOhHard] = b; S[kOhHardSum] = c;
        S[kOhMode] = ((long long)b < n_min) ? 1.0 : 0.0;      // 1 = top-k fallback (core/loss.py:17-18)
        S[kOhK] = (double)n_min; S[kOhPrefix] = 0.0; S[kOhRemain] = (double)n_min;
    }
}
// one radix pass (8 bits at `shift`) over the values whose higher bits equal the running prefix
__global__ void __launch_bounds__(256) ohem_hist_kernel(const float* __restrict__ loss_pix, int64_t count, int shift,
                                                        const double* __restrict__ ws, unsigned int* __restrict__ hist) {
    const double* S = ws + kWsScalars;
    if (S[kOhMode] == 0.0) return;
    __shared__ unsigned int h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const unsigned int prefix = (unsigned int)S[kOhPrefix];
    const unsigned int himask = shift == 24 ? 0u : (0xffffffffu << (shift + 8));
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const unsigned int o = f2ord(loss_pix[i]);
        if ((o & himask) == (prefix & himask)) atomicAdd(&h[(o >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (h[threadIdx.x]) atomicAdd(&hist[threadIdx.x], h[threadIdx.x]);
}
__global__ void ohem_scan_kernel(double* __restrict__ ws, int shift, unsigned int* __restrict__ hist) {
    double* S = ws + kWsScalars;
    if (threadIdx.x == 0 && S[kOhMode] != 0.0) {
        long long remain = (long long)S[kOhRemain];
        unsigned int prefix = (unsigned int)S[kOhPrefix];
        int b = 255;
        for (; b > 0; --b) {                 // walk from the largest bucket down
            if ((long long)hist[b] >= remain) break;
            remain -= hist[b];
        }
        prefix |= ((unsigned int)b) << shift;
        S[kOhPrefix] = (double)prefix; S[kOhRemain] = (double)remain;
        if (shift == 0) S[kOhV] = (double)ord2f(prefix);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
}
__global__ void __launch_bounds__(256) ohem_topk_sum_kernel(const float* __restrict__ loss_pix, int64_t count, double* __restrict__ ws) {
    const double* S = ws + kWsScalars;
    if (S[kOhMode] == 0.0) return;
    const float v = (float)S[kOhV];
    float sg = 0.f, cg = 0.f, ce = 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const float l = loss_pix[i];
        if (l > v) { sg += l; cg += 1.f; } else if (l == v) ce += 1.f;
    }
    __shared__ double sh[32];
    const double a = block_sum<double>((double)sg, sh), b = block_sum<double>((double)cg, sh), c = block_sum<double>((double)ce, sh);
    if (threadIdx.x == 0) { ws[blockIdx.x] = a; ws[kLossMaxBlocks / 2 + blockIdx.x] = b; ws[kWsLoss + blockIdx.x] = c; }
}
__global__ void ohem_finalize_kernel(double* __restrict__ ws, int npart, float* loss_out) {
    double* S = ws + kWsScalars;
    if (S[kOhMode] == 0.0) {
        if (threadIdx.x == 0) *loss_out = (float)(S[kOhHardSum] / S[kOhHard]);     // 0/0 -> NaN like torch.mean(empty)
        return;
    }
    __shared__ double sh[32];
    double a = 0, b = 0, c = 0;
    for (int i = threadIdx.x; i < npart; i += blockDim.x) { a += ws[i]; b += ws[kLossMaxBlocks / 2 + i]; c += ws[kWsLoss + i]; }
    a = block_sum<double>(a, sh); b = block_sum<double>(b, sh); c = block_sum<double>(c, sh);
    if (threadIdx.x == 0) {
        S[kOhSumGt] = a; S[kOhCntGt] = b; S[kOhCntEq] = c;
        const double k = S[kOhK];
        *loss_out = (float)((a + (k - b) * S[kOhV]) / k);
    }
}
// per-pixel upstream coefficient for the CE gradient
__global__ void __launch_bounds__(256) ohem_coef_kernel(float* __restrict__ loss_pix, int64_t count, float th, const double* __restrict__ ws) {
    const double* S = ws + kWsScalars;
    const bool topk = S[kOhMode] != 0.0;
    const float v = (float)S[kOhV];
    const float inv = topk ? (float)(1.0 / S[kOhK]) : (float)(1.0 / S[kOhHard]);
    const float tie = topk && S[kOhCntEq] > 0.0 ? (float)((S[kOhK] - S[kOhCntGt]) / (S[kOhCntEq] * S[kOhK])) : 0.f;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (int64_t)gridDim.x * blockDim.x) {
        const float l = loss_pix[i];
        float cf;
        if (!topk) cf = l > th ? inv : 0.f;
        else cf = l > v ? inv : (l == v ? tie : 0.f);
        loss_pix[i] = cf;
    }
}

}  // namespace
}  // namespace msx

using namespace msx;

extern "C" {

size_t msx_ws_bytes_loss(void) { return (size_t)kWsDoubles * sizeof(double) + 256 * sizeof(unsigned int); }
static size_t ce_codes_offset() { return (msx_ws_bytes_loss() + 255) & ~(size_t)255; }
size_t msx_ws_bytes_ce(int64_t pixels) { return ce_codes_offset() + (size_t)((pixels + 15) & ~(int64_t)15); }

int msx_ce_loss_fwd_bwd(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw,
                        const float* class_weights, int64_t ignore_index, int reduction, float* loss_out, float* dlogits,
                        void* ws, size_t ws_bytes, void* stream) {
    MSX_REQUIRE(logits && labels && loss_out && ws, "ce_loss: null pointer");
    MSX_REQUIRE(n > 0 && c > 0 && hw > 0, "ce_loss: empty input");
    MSX_REQUIRE(reduction == MSX_REDUCE_MEAN || reduction == MSX_REDUCE_SUM, "ce_loss: use msx_ce_loss_none_* for reduction='none'");
    if (ws_bytes < msx_ws_bytes_loss()) return set_error(MSX_ERR_WORKSPACE, "ce_loss: workspace %zu < %zu", ws_bytes, msx_ws_bytes_loss());
    cudaStream_t st = as_stream(stream);
    double* w = static_cast<double*>(ws);
    const int use_denom = reduction == MSX_REDUCE_MEAN;
    int nw = 0, nl = 0;
    uint8_t* codes = nullptr;
    if (use_denom) {
        // With the larger msx_ws_bytes_ce() workspace the pre-pass also leaves one-byte label codes behind.
        if (c <= 255 && ws_bytes >= msx_ws_bytes_ce((int64_t)n * hw)) codes = static_cast<uint8_t*>(ws) + ce_codes_offset();
        const bool v4 = hw % 4 == 0;
        const int64_t groups = (int64_t)n * hw / (v4 ? 4 : 1);
        nw = loss_grid(groups);
        if (v4) { ProfScope ps__(K_LOSS, st, 0, 0); ce_wsum_kernel<4><<<nw, 256, 0, st>>>(labels, groups, c, class_weights, ignore_index, w, codes); }
        else { ProfScope ps__(K_LOSS, st, 0, 0); ce_wsum_kernel<1><<<nw, 256, 0, st>>>(labels, groups, c, class_weights, ignore_index, w, codes); }
        MSX_LAUNCH_CHECK();
    }
    MSX_TRY(ce_dispatch<0>(logits, labels, n, c, hw, class_weights, ignore_index, nw, use_denom, nullptr, nullptr, dlogits, w, &nl, st, codes));
    { ProfScope ps__(K_MISC, st, 0, 0); ce_finalize_kernel<<<1, 256, 0, st>>>(w, nw, nl, use_denom, loss_out); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_ce_loss_none_fwd(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw,
                         const float* class_weights, int64_t ignore_index, float* loss_pix, void* stream) {
    MSX_REQUIRE(logits && labels && loss_pix, "ce_loss_none_fwd: null pointer");
    int g;
    return ce_dispatch<1>(logits, labels, n, c, hw, class_weights, ignore_index, 0, 0, nullptr, loss_pix, nullptr, nullptr, &g, as_stream(stream));
}

int msx_ce_loss_none_bwd(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw,
                         const float* class_weights, int64_t ignore_index, const float* coef_pix, float* dlogits, void* stream) {
    MSX_REQUIRE(logits && labels && coef_pix && dlogits, "ce_loss_none_bwd: null pointer");
    int g;
    return ce_dispatch<2>(logits, labels, n, c, hw, class_weights, ignore_index, 0, 0, coef_pix, nullptr, dlogits, nullptr, &g, as_stream(stream));
}

int msx_ohem_loss_fwd_bwd(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw, float thresh_prob,
                          int64_t ignore_index, float* loss_out, float* dlogits, float* loss_pix, void* ws, size_t ws_bytes,
                          void* stream) {
    MSX_REQUIRE(logits && labels && loss_out && loss_pix && ws, "ohem: null pointer");
    if (ws_bytes < msx_ws_bytes_loss()) return set_error(MSX_ERR_WORKSPACE, "ohem: workspace too small");
    cudaStream_t st = as_stream(stream);
    double* w = static_cast<double*>(ws);
    unsigned int* hist = reinterpret_cast<unsigned int*>(w + kWsDoubles);
    const float th = -logf(thresh_prob);
    const int64_t count = (int64_t)n * hw;
    int g;
    MSX_TRY(ce_dispatch<1>(logits, labels, n, c, hw, nullptr, ignore_index, 0, 0, nullptr, loss_pix, nullptr, nullptr, &g, st));
    int grid = loss_grid(count);
    if (grid > kLossMaxBlocks / 2) grid = kLossMaxBlocks / 2;
    { ProfScope ps__(K_LOSS, st, (double)count * 12.0, 0); ohem_count_kernel<<<grid, 256, 0, st>>>(loss_pix, labels, count, ignore_index, th, w); }
    { ProfScope ps__(K_MISC, st, 0, 0); ohem_decide_kernel<<<1, 256, 0, st>>>(w, grid, hist); }
    for (int shift = 24; shift >= 0; shift -= 8) {
        { ProfScope ps__(K_LOSS, st, (double)count * 4.0, 0); ohem_hist_kernel<<<grid, 256, 0, st>>>(loss_pix, count, shift, w, hist); }
        { ProfScope ps__(K_MISC, st, 0, 0); ohem_scan_kernel<<<1, 256, 0, st>>>(w, shift, hist); }
    }
    { ProfScope ps__(K_LOSS, st, (double)count * 4.0, 0); ohem_topk_sum_kernel<<<grid, 256, 0, st>>>(loss_pix, count, w); }
    { ProfScope ps__(K_MISC, st, 0, 0); ohem_finalize_kernel<<<1, 256, 0, st>>>(w, grid, loss_out); }
    MSX_LAUNCH_CHECK();
    if (dlogits) {
        { ProfScope ps__(K_LOSS, st, (double)count * 8.0, 0); ohem_coef_kernel<<<grid, 256, 0, st>>>(loss_pix, count, th, w); }
        MSX_LAUNCH_CHECK();
        MSX_TRY(ce_dispatch<2>(logits, labels, n, c, hw, nullptr, ignore_index, 0, 0, loss_pix, nullptr, dlogits, nullptr, &g, st));
    }
    return MSX_OK;
}

int msx_kd_loss_fwd_bwd(const float* student, const float* teacher, int32_t n, int32_t c, int64_t hw, int kind,
                        float temperature, float* loss_out, float* dstudent, void* ws, size_t ws_bytes, void* stream) {
    MSX_REQUIRE(student && teacher && loss_out && ws, "kd_loss: null pointer");
    if (ws_bytes < msx_ws_bytes_loss()) return set_error(MSX_ERR_WORKSPACE, "kd_loss: workspace too small");
    cudaStream_t st = as_stream(stream);
    double* w = static_cast<double*>(ws);
    const double total = (double)n * c * (double)hw;
    int g;
    if (kind == 0) {
        MSX_REQUIRE(temperature > 0.f, "kd_loss: temperature must be positive");
        const bool v8 = hw % 8 == 0 && c <= 4, v4 = hw % 4 == 0;
        const int64_t groups = (int64_t)n * hw / (v8 ? 8 : (v4 ? 4 : 1));
        MSX_REQUIRE(groups_ok(groups), "kd_loss: more than 2^31 pixel groups");
        g = 1;
        const float gscale = (float)((double)temperature / total);      // T^2/(NCHW) * 1/T
        {
            ProfScope ps__(K_LOSS, st, (double)total * (dstudent ? 12.0 : 8.0), 0);
#define MSX_KD(PV, CT) g = resident_grid((const void*)kd_kl_reg_kernel<PV, CT>, 0, groups), kd_kl_reg_kernel<PV, CT><<<g, 256, 0, st>>>(student, teacher, c, hw, groups, 1.f / temperature, gscale, dstudent, w)
            if (v8) {
                if (c <= 2) MSX_KD(8, 2); else MSX_KD(8, 4);
            } else if (v4) {
                if (c <= 2) MSX_KD(4, 2); else if (c <= 4) MSX_KD(4, 4); else if (c <= 8) MSX_KD(4, 8);
                else g = resident_grid((const void*)kd_kl_kernel<4>, 0, groups), kd_kl_kernel<4><<<g, 256, 0, st>>>(student, teacher, c, hw, groups, 1.f / temperature, gscale, dstudent, w);
            } else {
                if (c <= 2) MSX_KD(1, 2); else if (c <= 4) MSX_KD(1, 4); else if (c <= 8) MSX_KD(1, 8);
                else g = resident_grid((const void*)kd_kl_kernel<1>, 0, groups), kd_kl_kernel<1><<<g, 256, 0, st>>>(student, teacher, c, hw, groups, 1.f / temperature, gscale, dstudent, w);
            }
#undef MSX_KD
        }
        { ProfScope ps__(K_MISC, st, 0, 0); scaled_sum_finalize_kernel<<<1, 256, 0, st>>>(w, g, (double)temperature * temperature / total, loss_out); }
    } else if (kind == 1) {
        const int64_t count = (int64_t)total;
        g = loss_grid(count);
        { ProfScope ps__(K_LOSS, st, (double)total * (dstudent ? 12.0 : 8.0), 0); kd_mse_kernel<<<g, 256, 0, st>>>(student, teacher, count, (float)(2.0 / total), dstudent, w); }
        { ProfScope ps__(K_MISC, st, 0, 0); scaled_sum_finalize_kernel<<<1, 256, 0, st>>>(w, g, 1.0 / total, loss_out); }
    } else {
        return set_error(MSX_ERR_UNSUPPORTED, "kd_loss: unknown kind %d", kind);
    }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_dice_bce_loss_fwd_bwd(const float* logits, const void* targets, int labels_are_float, int64_t pixels, float eps,
                              float* loss_out, float* dlogits, void* ws, size_t ws_bytes, void* stream) {
    MSX_REQUIRE(logits && targets && loss_out && ws && pixels > 0, "dice_bce: null pointer / empty");
    if (ws_bytes < msx_ws_bytes_loss()) return set_error(MSX_ERR_WORKSPACE, "dice_bce: workspace too small");
    cudaStream_t st = as_stream(stream);
    double* w = static_cast<double*>(ws);
    int g = labels_are_float ? resident_grid((const void*)dice_bce_reduce_kernel<true>, 0, (pixels + 3) / 4)
                             : resident_grid((const void*)dice_bce_reduce_kernel<false>, 0, (pixels + 3) / 4);
    if (g > kLossMaxBlocks / 2) g = kLossMaxBlocks / 2;
    if (labels_are_float) { ProfScope ps__(K_LOSS, st, (double)pixels * (labels_are_float ? 8.0 : 12.0), 0); dice_bce_reduce_kernel<true><<<g, 256, 0, st>>>(logits, targets, pixels, w); }
    else { ProfScope ps__(K_LOSS, st, (double)pixels * (labels_are_float ? 8.0 : 12.0), 0); dice_bce_reduce_kernel<false><<<g, 256, 0, st>>>(logits, targets, pixels, w); }
    { ProfScope ps__(K_MISC, st, 0, 0); dice_bce_finalize_kernel<<<1, 256, 0, st>>>(w, g, pixels, eps, loss_out); }
    if (dlogits) {
        const int g2 = labels_are_float ? resident_grid((const void*)dice_bce_grad_kernel<true>, 0, (pixels + 3) / 4)
                                        : resident_grid((const void*)dice_bce_grad_kernel<false>, 0, (pixels + 3) / 4);
        if (labels_are_float) { ProfScope ps__(K_LOSS, st, (double)pixels * (labels_are_float ? 12.0 : 16.0), 0); dice_bce_grad_kernel<true><<<g2, 256, 0, st>>>(logits, targets, pixels, w, dlogits); }
        else { ProfScope ps__(K_LOSS, st, (double)pixels * (labels_are_float ? 12.0 : 16.0), 0); dice_bce_grad_kernel<false><<<g2, 256, 0, st>>>(logits, targets, pixels, w, dlogits); }
    }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_scale_by_device_scalar(float* data, int64_t count, const float* scale_dev, void* stream) {
    MSX_REQUIRE(data && scale_dev, "scale: null pointer");
    { ProfScope ps__(K_LOSS, as_stream(stream), 0, 0); scale_kernel<<<loss_grid(count), 256, 0, as_stream(stream)>>>(data, count, scale_dev); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int msx_confmat_update(const float* logits, const int64_t* labels, int32_t n, int32_t c, int64_t hw, int64_t ignore_index,
                       int64_t* confmat, int32_t* bad_label_flag, void* stream) {
    MSX_REQUIRE(logits && labels && confmat && bad_label_flag, "confmat: null pointer");
    MSX_REQUIRE(c >= 1 && c <= 64, "confmat: num_classes %d not in [1,64]", c);
    const bool v4 = hw % 4 == 0;
    const int64_t groups = (int64_t)n * hw / (v4 ? 4 : 1);
    const size_t smem = (c <= 4 ? (size_t)c * c * 256 : (size_t)c * c) * sizeof(unsigned int);
    const int g = v4 ? resident_grid((const void*)confmat_kernel<4>, smem, groups) : resident_grid((const void*)confmat_kernel<1>, smem, groups);
    auto* cm = reinterpret_cast<unsigned long long*>(confmat);
    if (v4) { ProfScope ps__(K_CONFMAT, as_stream(stream), (double)n * hw * (4.0 * c + 8.0), 0); confmat_kernel<4><<<g, 256, smem, as_stream(stream)>>>(logits, labels, c, hw, groups, ignore_index, cm, bad_label_flag); }
    else { ProfScope ps__(K_CONFMAT, as_stream(stream), (double)n * hw * (4.0 * c + 8.0), 0); confmat_kernel<1><<<g, 256, smem, as_stream(stream)>>>(logits, labels, c, hw, groups, ignore_index, cm, bad_label_flag); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

}  // extern "C"
