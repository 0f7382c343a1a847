// SIMT (FFMA, fp32-accumulate) implicit-GEMM convolution engine: the true-fp32 path of the
// library (parity mode: the reference's CPU convolutions are exact fp32) and
// the engine for shapes the tcgen05 kernel does not cover.  Handles every geometry the reference
// instantiates - 3x3 (d1/d2/d3), 3x3 s2, 2x2 s2, 1x1, 1x7, 7x1 and ConvTranspose2d 3x3 s2 -
// through one "gather" kernel (fprop / dgrad / transposed-conv forward) and one "wgrad" kernel.
//
// Gather form:  dst[p][n] = sum_{tap,k} src[ map(p,tap) ][k] * Wp[tap][n][k]
//   F-form (Conv2d fprop, ConvTranspose2d dgrad):  sy = oy*stride + r*dil - pad
//   B-form (Conv2d dgrad, ConvTranspose2d fprop):  sy = (oy + pad - r*dil) / stride  (if divisible)
// Wp is the packed operand written by msx_conv_pack_weights: [tap][Np=round16(N)][Kp=round8(K)],
// zero padded, in the activation dtype.
#include "common.cuh"

#include <vector>

namespace msx {

struct GatherGeom {
    int N, Hs, Ws, Hd, Wd;  // batch, source and destination spatial size
    int K, Nc;              // source channels (reduction) and destination channels
    int Kp, Np;             // packed operand dims
    int lds, ldd;
    int kh, kw;
    int mul_d, mul_t, off_h, off_w, den;
};

struct WgradGeom {
    int N, Hs, Ws;   // "small" side spatial (the F-form destination)
    int Hb, Wb;      // "big" side spatial (the F-form source)
    int A, B;        // channels of small / big;  dW layout [A][B][taps]
    int lda, ldb;
    int kh, kw, stride, dil, pad_h, pad_w;
    int tiles_a, tiles_b, splits;
    int64_t chunk;   // pixels per split
};

namespace {

__device__ inline bool map_coord(int o, int tap, int mul_d, int mul_t, int off, int den, int lim, int& s) {
    int v = o * mul_d + tap * mul_t + off;
    if (v < 0) return false;
    if (den > 1) {
        if (v % den) return false;
        v /= den;
    }
    s = v;
    return v < lim;
}

template <typename T, int TN>
__global__ void __launch_bounds__(256) conv_gather_kernel(const T* __restrict__ src, const T* __restrict__ wp,
                                                          const ConvEpilogue ep, T* __restrict__ dst, GatherGeom g) {
    constexpr int BM = 64, BK = 16, BN = 16 * TN;
    __shared__ __align__(16) float As[BK][BM + 4];
    __shared__ __align__(16) float Bs[BK][BN + 4];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int64_t M = (int64_t)g.N * g.Hd * g.Wd;
    const int64_t m0 = (int64_t)blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;

    // load roles: pixel / weight row (tid >> 2), 4-channel group (tid & 3)
    const int lrow = tid >> 2, kq = (tid & 3) * 4;
    const int64_t lp = m0 + lrow;
    const bool pvalid = lp < M;
    int pn = 0, poy = 0, pox = 0;
    if (pvalid) {
        pox = (int)(lp % g.Wd);
        const int64_t r = lp / g.Wd;
        poy = (int)(r % g.Hd);
        pn = (int)(r / g.Hd);
    }
    const bool brow = lrow < BN && (n0 + lrow) < g.Np;

    float acc[4][TN];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < TN; ++j) acc[i][j] = 0.f;

    for (int r = 0; r < g.kh; ++r) {
        for (int s = 0; s < g.kw; ++s) {
            int sy = 0, sx = 0;
            const bool ok = pvalid && map_coord(poy, r, g.mul_d, g.mul_t, g.off_h, g.den, g.Hs, sy) &&
                            map_coord(pox, s, g.mul_d, g.mul_t, g.off_w, g.den, g.Ws, sx);
            const T* sp = src + ((int64_t)(pn * g.Hs + sy) * g.Ws + sx) * g.lds;
            const T* bp = wp + ((int64_t)(r * g.kw + s) * g.Np + n0 + lrow) * g.Kp;
            for (int k0 = 0; k0 < g.Kp; k0 += BK) {
                float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
                const bool kin = (k0 + kq) < g.Kp;
                if (ok && kin) load4(sp + k0 + kq, a);
                if (brow && kin) load4(bp + k0 + kq, b);
                __syncthreads();
#pragma unroll
                for (int i = 0; i < 4; ++i) As[kq + i][lrow] = a[i];
                if (lrow < BN) {
#pragma unroll
                    for (int i = 0; i < 4; ++i) Bs[kq + i][lrow] = b[i];
                }
                __syncthreads();
#pragma unroll
                for (int kk = 0; kk < BK; ++kk) {
                    const float4 av = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
                    const float am[4] = {av.x, av.y, av.z, av.w};
                    float bn[TN];
                    if constexpr (TN == 4) {
                        const float4 bv = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
                        bn[0] = bv.x; bn[1] = bv.y; bn[2] = bv.z; bn[3] = bv.w;
                    } else {
                        const float2 bv = *reinterpret_cast<const float2*>(&Bs[kk][tx * 2]);
                        bn[0] = bv.x; bn[1] = bv.y;
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < TN; ++j) acc[i][j] = fmaf(am[i], bn[j], acc[i][j]);
                }
            }
        }
    }
    const int ncp = round_up(g.Nc, 8);
    const int nb = n0 + tx * TN;
    if (nb >= ncp) return;
    float bv[TN], sc[TN], sh[TN];
#pragma unroll
    for (int j = 0; j < TN; ++j) {
        const bool in = nb + j < g.Nc;
        bv[j] = (ep.bias && in) ? ep.bias[nb + j] : 0.f;
        sc[j] = ep.scale ? (in ? ep.scale[nb + j] : 0.f) : 1.f;
        sh[j] = (ep.scale && in) ? ep.shift[nb + j] : 0.f;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int64_t p = m0 + ty * 4 + i;
        if (p >= M) continue;
        float o[TN];
#pragma unroll
        for (int j = 0; j < TN; ++j) {
            o[j] = acc[i][j] + bv[j];
            if (ep.scale) o[j] = fmaf(o[j], sc[j], sh[j]);
            if (ep.relu) o[j] = fmaxf(o[j], 0.f);
        }
        if constexpr (TN == 4) store4(dst + p * g.ldd + nb, o);
        else store2(dst + p * g.ldd + nb, o[0], o[1]);
    }
}

// dW partial for one (tile_a, tile_b, tap, split):  part[a][b] = sum_p small[p][a] * big[F(p,tap)][b]
template <typename T, int TT>
__global__ void __launch_bounds__(256) conv_wgrad_kernel(const T* __restrict__ sm, const T* __restrict__ bg,
                                                         float* __restrict__ partial, WgradGeom g) {
    constexpr int BT = 16 * TT, BK = 16;
    __shared__ __align__(16) float As[BK][BT + 4];
    __shared__ __align__(16) float Bs[BK][BT + 4];
    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int ta = blockIdx.x / g.tiles_b, tb = blockIdx.x % g.tiles_b;
    const int tap = blockIdx.y, r = tap / g.kw, s = tap % g.kw;
    const int a0 = ta * BT, b0 = tb * BT;
    const int64_t M = (int64_t)g.N * g.Hs * g.Ws;
    const int64_t pbeg = (int64_t)blockIdx.z * g.chunk;
    const int64_t pend = pbeg + g.chunk < M ? pbeg + g.chunk : M;
    const int acp = round_up(g.A, 8), bcp = round_up(g.B, 8);

    constexpr int VPR = BT / 4;                 // float4 per pixel row of the tile
    const int lpx = tid / VPR, lc = (tid % VPR) * 4;
    const bool loader = tid < BK * VPR;

    float acc[TT][TT];
#pragma unroll
    for (int i = 0; i < TT; ++i)
#pragma unroll
        for (int j = 0; j < TT; ++j) acc[i][j] = 0.f;

    for (int64_t p0 = pbeg; p0 < pend; p0 += BK) {
        float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
        if (loader) {
            const int64_t p = p0 + lpx;
            if (p < pend) {
                if (a0 + lc < acp) load4(sm + p * g.lda + a0 + lc, a);
                const int ox = (int)(p % g.Ws);
                const int64_t q = p / g.Ws;
                const int oy = (int)(q % g.Hs);
                const int nn = (int)(q / g.Hs);
                const int sy = oy * g.stride + r * g.dil - g.pad_h;
                const int sx = ox * g.stride + s * g.dil - g.pad_w;
                if (sy >= 0 && sy < g.Hb && sx >= 0 && sx < g.Wb && b0 + lc < bcp)
                    load4(bg + ((int64_t)(nn * g.Hb + sy) * g.Wb + sx) * g.ldb + b0 + lc, b);
            }
        }
        __syncthreads();
        if (loader) {
            *reinterpret_cast<float4*>(&As[lpx][lc]) = make_float4(a[0], a[1], a[2], a[3]);
            *reinterpret_cast<float4*>(&Bs[lpx][lc]) = make_float4(b[0], b[1], b[2], b[3]);
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float am[TT], bn[TT];
#pragma unroll
            for (int i = 0; i < TT; ++i) { am[i] = As[kk][ty * TT + i]; bn[i] = Bs[kk][tx * TT + i]; }
#pragma unroll
            for (int i = 0; i < TT; ++i)
#pragma unroll
                for (int j = 0; j < TT; ++j) acc[i][j] = fmaf(am[i], bn[j], acc[i][j]);
        }
    }
    const int Apad = g.tiles_a * BT, Bpad = g.tiles_b * BT;
    float* out = partial + ((int64_t)(blockIdx.z * (g.kh * g.kw) + tap) * Apad) * Bpad;
#pragma unroll
    for (int i = 0; i < TT; ++i)
#pragma unroll
        for (int j = 0; j < TT; ++j) out[(int64_t)(a0 + ty * TT + i) * Bpad + b0 + tx * TT + j] = acc[i][j];
}

// dW[(a*B + b)*taps + tap] = sum_split partial[split][tap][a][b]   (reference OIHW / IOHW layout)
__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const float* __restrict__ partial, int splits, int taps, int A, int B,
                                                           int Apad, int Bpad, float* __restrict__ dw) {
    pdl_wait();
    pdl_trigger();
    // A CTA = 32 consecutive outputs x 8 slices of the split range.  Outputs are ordered (tap, a, b) with b
    // fastest so the partial reads are coalesced; each slice sums its splits in fixed order (4 loads in
    // flight), the 8 slice sums are combined in fixed order through shared memory: deterministic.
    __shared__ float sh[8][33];
    const int64_t total = (int64_t)A * B * taps;
    const int64_t zstride = (int64_t)taps * Apad * Bpad;
    const int el = threadIdx.x & 31, sub = threadIdx.x >> 5;
    const int per = (splits + 7) / 8;
    const int z0 = sub * per, z1 = min(splits, z0 + per);
    for (int64_t base = (int64_t)blockIdx.x * 32; base < total; base += (int64_t)gridDim.x * 32) {
        const int64_t i = base + el;
        float acc = 0.f;
        int a = 0, b = 0, tap = 0;
        if (i < total) {
            b = (int)(i % B);
            const int64_t q = i / B;
            a = (int)(q % A); tap = (int)(q / A);
            const float* src = partial + ((int64_t)tap * Apad + a) * Bpad + b;
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
            int z = z0;
            for (; z + 3 < z1; z += 4) {
                s0 += src[(z + 0) * zstride];
                s1 += src[(z + 1) * zstride];
                s2 += src[(z + 2) * zstride];
                s3 += src[(z + 3) * zstride];
            }
            for (; z < z1; ++z) s0 += src[z * zstride];
            acc = (s0 + s1) + (s2 + s3);
        }
        sh[sub][el] = acc;
        __syncthreads();
        if (sub == 0 && i < total) {
            float t = 0.f;
#pragma unroll
            for (int k = 0; k < 8; ++k) t += sh[k][el];
            dw[((int64_t)a * B + b) * taps + tap] = t;
        }
        __syncthreads();
    }
}

// Few splits (wide layers): one thread per output, reads coalesced along b.
__global__ void __launch_bounds__(256) wgrad_reduce_flat_kernel(const float* __restrict__ partial, int splits, int taps, int A, int B,
                                                                int Apad, int Bpad, float* __restrict__ dw) {
    pdl_wait();
    pdl_trigger();
    const int64_t total = (int64_t)A * B * taps;
    const int64_t zstride = (int64_t)taps * Apad * Bpad;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int b = (int)(i % B);
        const int64_t q = i / B;
        const int a = (int)(q % A), tap = (int)(q / A);
        const float* src = partial + ((int64_t)tap * Apad + a) * Bpad + b;
        float s = 0.f;
        for (int z = 0; z < splits; ++z) s += src[z * zstride];
        dw[((int64_t)a * B + b) * taps + tap] = s;
    }
}

// master W[A][B][taps] fp32 -> packed [tap][Np][Kp] (n = a or b)
template <typename T>
__global__ void pack_weights_kernel(const float* __restrict__ w, int A, int B, int taps, int n_is_a, int Np, int Kp,
                                    T* __restrict__ out) {
    const int64_t total = (int64_t)taps * Np * Kp;
    const int Nn = n_is_a ? A : B, Kk = n_is_a ? B : A;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int k = (int)(i % Kp);
        const int64_t q = i / Kp;
        const int n = (int)(q % Np), tap = (int)(q / Np);
        float v = 0.f;
        if (n < Nn && k < Kk) {
            const int a = n_is_a ? n : k, b = n_is_a ? k : n;
            v = w[((int64_t)a * B + b) * taps + tap];
        }
        out[i] = from_float<T>(v);
    }
}

// ---- all weights of a model in one launch ---------------------------------------------------
// A step packs every conv weight twice (fprop and dgrad operand).  Per-conv launches of the kernel above were
// ~540 launches / step for DUCK-Net; the table-driven kernel below does the same work in one.
// Work unit = one (n, k) pair of one item: the thread reads that pair's `taps` master values (contiguous when
// k runs along the master's inner axis) and writes one element of every tap plane.
struct PackItemDev {
    const float* w;
    void* out;
    int A, B, taps, n_is_a, Np, Kp;
};
struct PackChunkDev { int item, first_pair; };
constexpr int kPackPairsPerCta = 1024;

template <typename T>
__global__ void __launch_bounds__(256) pack_many_kernel(const PackItemDev* __restrict__ items, const PackChunkDev* __restrict__ chunks) {
    const PackChunkDev ch = chunks[blockIdx.x];
    const PackItemDev it = items[ch.item];
    const int pairs = it.Np * it.Kp;
    const int Nn = it.n_is_a ? it.A : it.B, Kk = it.n_is_a ? it.B : it.A;
    T* out = static_cast<T*>(it.out);
    const int end = min(ch.first_pair + kPackPairsPerCta, pairs);
    for (int i = ch.first_pair + threadIdx.x; i < end; i += 256) {
        const int n = i / it.Kp, k = i - n * it.Kp;
        if (n < Nn && k < Kk) {
            const int a = it.n_is_a ? n : k, b = it.n_is_a ? k : n;
            const float* src = it.w + ((int64_t)a * it.B + b) * it.taps;
            for (int tap = 0; tap < it.taps; ++tap) out[(int64_t)tap * pairs + i] = from_float<T>(src[tap]);
        } else {
            for (int tap = 0; tap < it.taps; ++tap) out[(int64_t)tap * pairs + i] = from_float<T>(0.f);
        }
    }
}

}  // namespace

static void pack_item_dims(const msx_pack_item& h, PackItemDev& d) {
    const bool fwd_pass = h.mode == MSX_PACK_FPROP;
    const int Nc = fwd_pass ? h.cout : h.cin, K = fwd_pass ? h.cin : h.cout;
    d.w = h.w_master;
    d.out = h.packed;
    d.A = h.transposed ? h.cin : h.cout;
    d.B = h.transposed ? h.cout : h.cin;
    d.taps = h.kh * h.kw;
    d.n_is_a = (h.transposed ? !fwd_pass : fwd_pass) ? 1 : 0;
    d.Np = round_up(Nc, 16);
    d.Kp = round_up(K, 8);
}

static int pack_plan_chunks(const msx_pack_item* items, int count) {
    int64_t chunks = 0;
    for (int i = 0; i < count; ++i) {
        PackItemDev d;
        pack_item_dims(items[i], d);
        chunks += ((int64_t)d.Np * d.Kp + kPackPairsPerCta - 1) / kPackPairsPerCta;
    }
    return chunks > INT32_MAX ? -1 : (int)chunks;
}

size_t conv_pack_plan_bytes(const msx_pack_item* items, int count) {
    const int chunks = pack_plan_chunks(items, count);
    if (chunks < 0) return 0;
    return (size_t)count * sizeof(PackItemDev) + (size_t)chunks * sizeof(PackChunkDev) + 256;
}

// Writes the table through a synchronous copy: building a plan is start-up work, not part of a step.
int conv_pack_plan_build(const msx_pack_item* items, int count, void* table, size_t table_bytes, int* chunks_out) {
    const int chunks = pack_plan_chunks(items, count);
    if (chunks < 0) return set_error(MSX_ERR_UNSUPPORTED, "pack_plan: too many chunks");
    if (table_bytes < conv_pack_plan_bytes(items, count)) return set_error(MSX_ERR_WORKSPACE, "pack_plan: table %zu < %zu bytes", table_bytes, conv_pack_plan_bytes(items, count));
    std::vector<PackItemDev> hi(count);
    std::vector<PackChunkDev> hc;
    hc.reserve(chunks);
    for (int i = 0; i < count; ++i) {
        const msx_pack_item& h = items[i];
        if (!h.w_master || !h.packed || h.cout <= 0 || h.cin <= 0 || h.kh <= 0 || h.kw <= 0 || (h.mode != MSX_PACK_FPROP && h.mode != MSX_PACK_DGRAD))
            return set_error(MSX_ERR_INVALID, "pack_plan: bad item %d", i);
        pack_item_dims(h, hi[i]);
        const int pairs = hi[i].Np * hi[i].Kp;
        for (int f = 0; f < pairs; f += kPackPairsPerCta) hc.push_back({i, f});
    }
    char* base = static_cast<char*>(table);
    MSX_CHECK_CUDA(cudaMemcpy(base, hi.data(), hi.size() * sizeof(PackItemDev), cudaMemcpyHostToDevice));
    const size_t off = ((size_t)count * sizeof(PackItemDev) + 15) & ~(size_t)15;
    MSX_CHECK_CUDA(cudaMemcpy(base + off, hc.data(), hc.size() * sizeof(PackChunkDev), cudaMemcpyHostToDevice));
    *chunks_out = chunks;
    return MSX_OK;
}

int conv_pack_plan_run(int dtype, const void* table, int count, int chunks, double bytes, cudaStream_t st) {
    const char* base = static_cast<const char*>(table);
    const size_t off = ((size_t)count * sizeof(PackItemDev) + 15) & ~(size_t)15;
    MSX_DISPATCH_DTYPE(dtype, T, {
        ProfScope ps__(K_PACK, st, bytes, 0);
        pack_many_kernel<T><<<chunks, 256, 0, st>>>(reinterpret_cast<const PackItemDev*>(base), reinterpret_cast<const PackChunkDev*>(base + off));
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

// --------------------------------------------------------------------------------------------
int validate_desc(const msx_conv_desc* d) {
    if (!d) return set_error(MSX_ERR_INVALID, "conv: null descriptor");
    if (d->n <= 0 || d->h <= 0 || d->w <= 0 || d->cin <= 0 || d->cout <= 0 || d->kh <= 0 || d->kw <= 0 || d->stride <= 0 || d->dil <= 0)
        return set_error(MSX_ERR_INVALID, "conv: non-positive dimension");
    int eoh, eow;
    if (!d->transposed) {
        eoh = (d->h + 2 * d->pad_h - d->dil * (d->kh - 1) - 1) / d->stride + 1;
        eow = (d->w + 2 * d->pad_w - d->dil * (d->kw - 1) - 1) / d->stride + 1;
        if (d->oh != eoh || d->ow != eow) return set_error(MSX_ERR_INVALID, "conv: output %dx%d inconsistent with geometry (expect %dx%d)", d->oh, d->ow, eoh, eow);
    } else {
        // output_padding is implied by oh: must lie in [0, stride)
        const int base_h = (d->h - 1) * d->stride - 2 * d->pad_h + d->dil * (d->kh - 1) + 1;
        const int base_w = (d->w - 1) * d->stride - 2 * d->pad_w + d->dil * (d->kw - 1) + 1;
        if (d->oh < base_h || d->oh >= base_h + d->stride || d->ow < base_w || d->ow >= base_w + d->stride)
            return set_error(MSX_ERR_INVALID, "conv_transpose: output %dx%d inconsistent with geometry", d->oh, d->ow);
    }
    if (d->ldx % 8 || d->ldy % 8 || d->ldx < round_up(d->cin, 8) || d->ldy < round_up(d->cout, 8))
        return set_error(MSX_ERR_INVALID, "conv: ldx=%d / ldy=%d must be multiples of 8 and >= padded channels", d->ldx, d->ldy);
    if (d->dtype != MSX_F32 && d->dtype != MSX_BF16) return set_error(MSX_ERR_INVALID, "conv: bad dtype");
    return MSX_OK;
}

// Operand dims of a pass. fwd_pass: true for the module's forward.
static void pass_dims(const msx_conv_desc* d, bool fwd_pass, int& Nc, int& K) {
    Nc = fwd_pass ? d->cout : d->cin;
    K = fwd_pass ? d->cin : d->cout;
}

size_t conv_packed_elems(const msx_conv_desc* d) {
    // both passes fit: taps * round16(max) * round8(max) is an upper bound; use exact max of the two
    const int taps = d->kh * d->kw;
    const size_t f = (size_t)taps * round_up(d->cout, 16) * round_up(d->cin, 8);
    const size_t b = (size_t)taps * round_up(d->cin, 16) * round_up(d->cout, 8);
    return f > b ? f : b;
}

int conv_pack_launch(const msx_conv_desc* d, int mode, const float* w, void* packed, cudaStream_t st) {
    const bool fwd_pass = mode == MSX_PACK_FPROP;
    int Nc, K;
    pass_dims(d, fwd_pass, Nc, K);
    const int Np = round_up(Nc, 16), Kp = round_up(K, 8), taps = d->kh * d->kw;
    // master layout [A][B][taps]: Conv2d A=cout,B=cin ; ConvTranspose2d A=cin,B=cout
    const int A = d->transposed ? d->cin : d->cout, B = d->transposed ? d->cout : d->cin;
    // n index runs over A when (conv & fwd) or (convT & bwd)
    const int n_is_a = (d->transposed ? !fwd_pass : fwd_pass) ? 1 : 0;
    const int64_t total = (int64_t)taps * Np * Kp;
    const int grid = (int)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096);
    MSX_DISPATCH_DTYPE(d->dtype, T, {
        { ProfScope ps__(K_PACK, st, (double)total * (4 + sizeof(T)), 0); pack_weights_kernel<T><<<grid, 256, 0, st>>>(w, A, B, taps, n_is_a, Np, Kp, static_cast<T*>(packed)); }
    });
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

// fwd_pass: module forward (x->y) else data gradient (dy->dx).
int conv_gather_simt(const msx_conv_desc* d, bool fwd_pass, const void* src, const void* packed, const ConvEpilogue& ep,
                     void* dst, cudaStream_t st) {
    GatherGeom g;
    g.N = d->n;
    int Nc, K;
    pass_dims(d, fwd_pass, Nc, K);
    g.K = K; g.Nc = Nc; g.Kp = round_up(K, 8); g.Np = round_up(Nc, 16);
    g.kh = d->kh; g.kw = d->kw;
    if (fwd_pass) { g.Hs = d->h; g.Ws = d->w; g.Hd = d->oh; g.Wd = d->ow; g.lds = d->ldx; g.ldd = d->ldy; }
    else          { g.Hs = d->oh; g.Ws = d->ow; g.Hd = d->h; g.Wd = d->w; g.lds = d->ldy; g.ldd = d->ldx; }
    const bool f_form = d->transposed ? !fwd_pass : fwd_pass;
    if (f_form) { g.mul_d = d->stride; g.mul_t = d->dil; g.off_h = -d->pad_h; g.off_w = -d->pad_w; g.den = 1; }
    else        { g.mul_d = 1; g.mul_t = -d->dil; g.off_h = d->pad_h; g.off_w = d->pad_w; g.den = d->stride; }
    const int64_t M = (int64_t)g.N * g.Hd * g.Wd;
    const int ncp = round_up(Nc, 8);
    MSX_DISPATCH_DTYPE(d->dtype, T, {
        if (ncp <= 32) {
            dim3 grid((unsigned)cdiv64(M, 64), (ncp + 31) / 32);
            { ProfScope ps__(K_CONV_SIMT, st, (double)M * (g.K + g.Nc) * (double)sizeof(T), 2.0 * (double)M * g.Nc * g.K * g.kh * g.kw); conv_gather_kernel<T, 2><<<grid, 256, 0, st>>>(static_cast<const T*>(src), static_cast<const T*>(packed), ep, static_cast<T*>(dst), g); }
        } else {
            dim3 grid((unsigned)cdiv64(M, 64), (ncp + 63) / 64);
            { ProfScope ps__(K_CONV_SIMT, st, (double)M * (g.K + g.Nc) * (double)sizeof(T), 2.0 * (double)M * g.Nc * g.K * g.kh * g.kw); conv_gather_kernel<T, 4><<<grid, 256, 0, st>>>(static_cast<const T*>(src), static_cast<const T*>(packed), ep, static_cast<T*>(dst), g); }
        }
    });
    MSX_LAUNCH_CHECK();
    note_conv_engine(MSX_ENGINE_SIMT);
    return MSX_OK;
}

static void wgrad_plan(const msx_conv_desc* d, WgradGeom& g, int& TT) {
    // F-form destination ("small") and source ("big") sides
    if (!d->transposed) { g.Hs = d->oh; g.Ws = d->ow; g.Hb = d->h; g.Wb = d->w; g.A = d->cout; g.B = d->cin; g.lda = d->ldy; g.ldb = d->ldx; }
    else                { g.Hs = d->h; g.Ws = d->w; g.Hb = d->oh; g.Wb = d->ow; g.A = d->cin; g.B = d->cout; g.lda = d->ldx; g.ldb = d->ldy; }
    g.N = d->n; g.kh = d->kh; g.kw = d->kw; g.stride = d->stride; g.dil = d->dil; g.pad_h = d->pad_h; g.pad_w = d->pad_w;
    const int mx = g.A > g.B ? g.A : g.B;
    TT = mx <= 32 ? 2 : 4;
    const int BT = 16 * TT;
    g.tiles_a = (g.A + BT - 1) / BT;
    g.tiles_b = (g.B + BT - 1) / BT;
    const int64_t M = (int64_t)g.N * g.Hs * g.Ws;
    const int64_t base = (int64_t)g.tiles_a * g.tiles_b * g.kh * g.kw;
    int64_t splits = (kNumSMs * 4 + base - 1) / base;
    const int64_t max_splits = cdiv64(M, 256);
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    g.chunk = cdiv64(cdiv64(M, splits), 16) * 16;
    g.splits = (int)cdiv64(M, g.chunk);
}

size_t conv_wgrad_ws_bytes(const msx_conv_desc* d) {
    WgradGeom g;
    int TT;
    wgrad_plan(d, g, TT);
    const int BT = 16 * TT;
    return (size_t)g.splits * g.kh * g.kw * (g.tiles_a * BT) * (size_t)(g.tiles_b * BT) * sizeof(float);
}

// x: module input, dy: gradient of module output; dw in the reference layout.
int conv_wgrad_simt(const msx_conv_desc* d, const void* x, const void* dy, float* dw, void* ws, size_t ws_bytes,
                    cudaStream_t st) {
    WgradGeom g;
    int TT;
    wgrad_plan(d, g, TT);
    const size_t need = conv_wgrad_ws_bytes(d);
    if (ws_bytes < need) return set_error(MSX_ERR_WORKSPACE, "conv wgrad: workspace %zu < %zu", ws_bytes, need);
    const void* sm = d->transposed ? x : dy;
    const void* bg = d->transposed ? dy : x;
    const int taps = g.kh * g.kw, BT = 16 * TT;
    dim3 grid(g.tiles_a * g.tiles_b, taps, g.splits);
    MSX_DISPATCH_DTYPE(d->dtype, T, {
        if (TT == 2)
            { ProfScope ps__(K_WGRAD_SIMT, st, (double)g.N * ((double)g.Hs * g.Ws * g.A + (double)g.Hb * g.Wb * g.B) * (double)sizeof(T), 2.0 * (double)g.N * g.Hs * g.Ws * g.A * g.B * taps); conv_wgrad_kernel<T, 2><<<grid, 256, 0, st>>>(static_cast<const T*>(sm), static_cast<const T*>(bg), static_cast<float*>(ws), g); }
        else
            { ProfScope ps__(K_WGRAD_SIMT, st, (double)g.N * ((double)g.Hs * g.Ws * g.A + (double)g.Hb * g.Wb * g.B) * (double)sizeof(T), 2.0 * (double)g.N * g.Hs * g.Ws * g.A * g.B * taps); conv_wgrad_kernel<T, 4><<<grid, 256, 0, st>>>(static_cast<const T*>(sm), static_cast<const T*>(bg), static_cast<float*>(ws), g); }
    });
    MSX_LAUNCH_CHECK();
    const int64_t total = (int64_t)g.A * g.B * taps;
    const int rg = (int)((total + 31) / 32 < 4096 ? (total + 31) / 32 : 4096);
    { ProfScope ps__(K_WGRAD_REDUCE, st, (double)total * 4 * (g.splits + 1), 0); if (g.splits >= 16) launch_pdl(wgrad_reduce_kernel, dim3(rg), dim3(256), 0, st, static_cast<const float*>(ws), g.splits, taps, g.A, g.B, g.tiles_a * BT, g.tiles_b * BT, dw); else launch_pdl(wgrad_reduce_flat_kernel, dim3((int)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096)), dim3(256), 0, st, static_cast<const float*>(ws), g.splits, taps, g.A, g.B, g.tiles_a * BT, g.tiles_b * BT, dw); }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

int wgrad_reduce_launch(const float* partial, int splits, int taps, int A, int B, int Apad, int Bpad, float* dw, cudaStream_t st) {
    const int64_t total = (int64_t)A * B * taps;
    const int rg = (int)((total + 31) / 32 < 4096 ? (total + 31) / 32 : 4096);
    {
        ProfScope ps__(K_WGRAD_REDUCE, st, (double)total * 4 * (splits + 1), 0);
        if (splits >= 16) launch_pdl(wgrad_reduce_kernel, dim3(rg), dim3(256), 0, st, partial, splits, taps, A, B, Apad, Bpad, dw);
        else launch_pdl(wgrad_reduce_flat_kernel, dim3((int)((total + 255) / 256 < 4096 ? (total + 255) / 256 : 4096)), dim3(256), 0, st, partial, splits, taps, A, B, Apad, Bpad, dw);
    }
    MSX_LAUNCH_CHECK();
    return MSX_OK;
}

}  // namespace msx
