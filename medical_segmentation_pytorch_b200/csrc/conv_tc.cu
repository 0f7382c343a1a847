// tcgen05 / TMEM / TMA implicit-GEMM convolution engine for sm_100a (bf16 in, fp32 accumulate).
//
// One kernel serves every stride-1 "gather" pass of the reference's convolutions (Conv2d fprop and
// dgrad for 3x3 d1/d2/d3, 1x1, 1x7, 7x1):   dst[p][n] = sum_{tap,k} src[p + off(tap)][k] * Wp[tap][n][k]
//
//   GEMM view     M = 128 output pixels (a th x tw patch of one image), N = BN <= 256 output channels,
//                 K = taps x channels, walked tap by tap in chunks of 64 (or 32) channels.
//   A operand     NHWC activations through a 4-D TMA tensor map {C, W, H, N}; the box {chunk, tw, th, 1}
//                 at (k0, x0 + dx(tap), y0 + dy(tap), img) lands in shared memory as 128 pixel rows of
//                 128 B (64 B) with the 128B (64B) swizzle = the canonical K-major UMMA layout.  Image
//                 borders (zero padding), ragged tiles and the channel tail are all TMA out-of-bounds
//                 zero fill: no im2col buffer, no predication.
//   B operand     packed weights [tap][Np][Kp] through a 3-D tensor map, box {chunk, BN, 1}.
//   D accumulator TMEM, 128 lanes x BN fp32 columns, double buffered (2 x BN <= 512 columns) so the
//                 epilogue of tile i overlaps the MMAs of tile i+1.
//   Warp roles    warp 0: TMA producer (one lane) - warp 1: TMEM alloc + tcgen05.mma issue (one lane) -
//                 warps 2..5: epilogue (tcgen05.ld -> +bias -> bf16 -> 16 B global stores).
//   Pipelines     smem ring of S stages (full/empty mbarriers, tcgen05.commit frees a stage);
//                 TMEM ring of 2 accumulators (tmem_full / tmem_empty mbarriers).
//   Scheduling    persistent: grid = min(#tiles, 148), static round-robin over (pixel tile, n tile).
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <set>
#include <utility>

#include "common.cuh"

namespace msx {

namespace {

constexpr int kThreads = 192;
constexpr int kHaloThreads = 320;      // TMA warp + MMA warp + 2 x 4 epilogue warps
constexpr int kHaloThreads2 = 352;     // halo kernel: + a second MMA-issuing warp (warp 10)
constexpr int kMaxStages = 8;
constexpr int kMaxAcc = 8;
#ifndef MSX_TC_SMEM_KB
#define MSX_TC_SMEM_KB 200
#endif
constexpr uint32_t kSmemBudget = MSX_TC_SMEM_KB * 1024;   // pipeline stages + resident weights of one CTA

constexpr int kTapMax = 49;          // tap table capacity (7x7); wgrad / dgrad stay within kTapMaxTrain
constexpr int kTapMaxTrain = 16;
struct TapTable {          // one entry per (source phase, shift, weight tap) the pass visits
    int n;
    signed char map[kTapMax], dx[kTapMax], dy[kTapMax], wtap[kTapMax];
};
struct PhaseMaps { CUtensorMap m[4]; };   // source seen through up to 4 stride-2 phase views

struct TcParams {
    int N, H, W;          // tile-grid domain: the destination image, or one stride-2 phase of it
    int Hd, Wd;           // full destination image size
    int dmul, da, db;     // destination pixel of grid point (y, x) is (y*dmul + da, x*dmul + db)
    int tw, th, tiles_x, tiles_y;
    int n_tiles, BN;      // tiling of the output-channel axis
    TapTable taps;
    int kchunks, chunk;   // K chunks per tap, channels per chunk (64 or 32)
    int tail;             // 0, or the box width (16 | 32 channels) of the LAST chunk when the channel count leaves a short tail
    uint32_t a_bytes_tail, b_bytes_tail;
    int Kp;               // padded reduction channels (multiple of 8)
    int Nc, ncp, ldd;     // output channels, round_up(Nc, 8), destination pixel stride
    int stages;
    uint32_t a_bytes, b_bytes, a_stage, b_stage;   // TMA box bytes and smem stage pitch
    int tmem_cols;
    int total_tiles;
};

// ---- PTX wrappers -----------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// bias / eval-BN affine / ReLU on NV accumulator columns starting at logical column col0 (uniform branches)
template <int NV>
__device__ __forceinline__ void epilogue_cols(uint32_t (&v)[NV], int col0, int Nc, const ConvEpilogue& ep) {
    if (ep.bias) {
#pragma unroll
        for (int j = 0; j < NV; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + (col0 + j < Nc ? ep.bias[col0 + j] : 0.f));
    }
    if (ep.scale) {
#pragma unroll
        for (int j = 0; j < NV; ++j) {
            float f = col0 + j < Nc ? fmaf(__uint_as_float(v[j]), ep.scale[col0 + j], ep.shift[col0 + j]) : 0.f;
            if (ep.relu) f = fmaxf(f, 0.f);
            v[j] = __float_as_uint(f);
        }
    } else if (ep.relu) {
#pragma unroll
        for (int j = 0; j < NV; ++j) v[j] = __float_as_uint(fmaxf(__uint_as_float(v[j]), 0.f));
    }
}

__device__ __forceinline__ uint32_t pack_bf16(uint32_t a, uint32_t b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(a), __uint_as_float(b));
    return *reinterpret_cast<uint32_t*>(&h);
}

// Column sums over the 32 lanes of a warp, one accumulator row per lane: on return lane l holds the sum over all lanes of
// v[l] (NV == 32) or of v[l & 15] (NV == 16).  Transpose-reduce: each step halves the columns a lane is responsible for,
// 31 shuffles in total instead of 5 per column.  The arrays are destroyed; element 0 holds the result.
template <int N>
__device__ __forceinline__ void colsum_step(float* a, float* b, int lane) {      // two independent reductions interleaved (ILP)
    const bool up = (lane & N) != 0;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const float ka = up ? a[i + N] : a[i], sa = up ? a[i] : a[i + N];
        const float kb = up ? b[i + N] : b[i], sb = up ? b[i] : b[i + N];
        a[i] = ka + __shfl_xor_sync(0xffffffffu, sa, N);
        b[i] = kb + __shfl_xor_sync(0xffffffffu, sb, N);
    }
}
template <int NV>
__device__ __forceinline__ void warp_colsum2(float (&a)[NV], float (&b)[NV], int lane) {
    if constexpr (NV == 32) colsum_step<16>(a, b, lane);
    else {
#pragma unroll
        for (int i = 0; i < 16; ++i) { a[i] += __shfl_xor_sync(0xffffffffu, a[i], 16); b[i] += __shfl_xor_sync(0xffffffffu, b[i], 16); }
    }
    colsum_step<8>(a, b, lane);
    colsum_step<4>(a, b, lane);
    colsum_step<2>(a, b, lane);
    colsum_step<1>(a, b, lane);
}

// One chunk of NV accumulator columns of one pixel row (this thread's): bias / affine / ReLU, bf16 pack, 16 B stores, and -
// with ep.stats - the training-mode BatchNorm statistics of exactly what was stored (the bf16-rounded values, so that the
// apply and backward passes, which re-read y, see the same numbers): per-column sums over the warp's 32 pixels in fp32,
// then a running double per (warp, column) in shared memory (stat_s / stat_q point at this chunk's column 0).
template <int NV>
__device__ __forceinline__ void epilogue_chunk(uint32_t (&v)[NV], int col0, int cols_left, bool valid, bf16* out, int Nc,
                                               const ConvEpilogue& ep, double* stat_s, double* stat_q, int lane) {
    epilogue_cols<NV>(v, col0, Nc, ep);
    uint32_t pk[NV / 2];
#pragma unroll
    for (int j = 0; j < NV / 2; ++j) pk[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
    if (valid) {
#pragma unroll
        for (int q = 0; q < NV / 8; ++q)
            if (8 * q < cols_left)
                *reinterpret_cast<uint4*>(out + 8 * q) = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
    }
    if (ep.stats) {
        float a[NV], b[NV];
#pragma unroll
        for (int j = 0; j < NV / 2; ++j) {
            a[2 * j] = valid ? __uint_as_float(pk[j] << 16) : 0.f;
            a[2 * j + 1] = valid ? __uint_as_float(pk[j] & 0xffff0000u) : 0.f;
            b[2 * j] = a[2 * j] * a[2 * j];
            b[2 * j + 1] = a[2 * j + 1] * a[2 * j + 1];
        }
        warp_colsum2<NV>(a, b, lane);
        if (NV == 32 || lane < 16) {
            stat_s[lane] += (double)a[0];
            stat_q[lane] += (double)b[0];
        }
    }
}

// The same chunk with the statistics DEFERRED: the thread adds its own pixel's values (and squares) to running fp32 sums in
// registers - it owns the same pixel position of every tile it drains, ~50 tiles per launch - and the 32-lane column reduction
// runs once per launch (stats_flush) instead of once per tile.  Measured before: the per-tile butterfly made the 17-channel
// forward convs epilogue bound, 66 -> 89 us with statistics.
template <int NV>
__device__ __forceinline__ void epilogue_chunk_defer(uint32_t (&v)[NV], int col0, int cols_left, bool valid, bf16* out, int Nc,
                                                     const ConvEpilogue& ep, float (&ra)[NV], float (&rb)[NV]) {
    epilogue_cols<NV>(v, col0, Nc, ep);
    uint32_t pk[NV / 2];
#pragma unroll
    for (int j = 0; j < NV / 2; ++j) pk[j] = pack_bf16(v[2 * j], v[2 * j + 1]);
    if (valid) {
#pragma unroll
        for (int q = 0; q < NV / 8; ++q)
            if (8 * q < cols_left)
                *reinterpret_cast<uint4*>(out + 8 * q) = make_uint4(pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
#pragma unroll
        for (int j = 0; j < NV / 2; ++j) {
            const float x0 = __uint_as_float(pk[j] << 16), x1 = __uint_as_float(pk[j] & 0xffff0000u);
            ra[2 * j] += x0; ra[2 * j + 1] += x1;
            rb[2 * j] = fmaf(x0, x0, rb[2 * j]); rb[2 * j + 1] = fmaf(x1, x1, rb[2 * j + 1]);
        }
    }
}
template <int NV>
__device__ __forceinline__ void stats_flush(float (&ra)[NV], float (&rb)[NV], double* stat_s, double* stat_q, int lane) {
    warp_colsum2<NV>(ra, rb, lane);
    if (NV == 32 || lane < 16) {
        stat_s[lane] += (double)ra[0];
        stat_q[lane] += (double)rb[0];
    }
}

// After every role of the CTA is done: fold the 8 epilogue warps' running sums into this CTA's row of partials (a finalize
// launch folds the <= 148 rows; with a ticket the last CTA of this launch does it instead).  All threads call.
__device__ __forceinline__ void stats_tail(const ConvEpilogue& ep, double* sstat, int nca, int* flag) {
    const BnFinalize& f = ep.bnf;
    for (int c = threadIdx.x; c < f.cp; c += blockDim.x) {
        double s = 0.0, q = 0.0;
#pragma unroll
        for (int w = 0; w < 8; ++w) { s += sstat[(w * 2 + 0) * nca + c]; q += sstat[(w * 2 + 1) * nca + c]; }
        f.partials[((int64_t)blockIdx.x * 2 + 0) * f.cp + c] = s;
        f.partials[((int64_t)blockIdx.x * 2 + 1) * f.cp + c] = q;
    }
    if (f.ticket && ticket_is_last(f.ticket, gridDim.x, flag)) bn_finalize_all(f, gridDim.x);
}

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
    return ok != 0;
}
// Bounded spin: a protocol bug traps (reported as a CUDA error) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t spins = 0;
    while (!mbar_try_wait(bar, parity)) {
        if (++spins > 50000000u) __trap();
    }
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// Same instruction with the descriptors assembled from 32-bit halves inside the asm block and a compile-time
// accumulate flag: the single issuing thread is bound by its own instruction latency, so every integer
// instruction per MMA counts (no "memory" clobber: asm volatile statements keep their mutual order).
template <int ACC>
__device__ __forceinline__ void umma_bf16_lh(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc) {
    asm volatile(
        "{\n"
        ".reg .b64 da, db;\n"
        ".reg .pred p;\n"
        "mov.b64 da, {%1, %2};\n"
        "mov.b64 db, {%3, %4};\n"
        "setp.ne.b32 p, %6, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n"
        "}\n" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "n"(ACC));
}
__device__ __forceinline__ void umma_bf16_lhr(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n"
        ".reg .b64 da, db;\n"
        ".reg .pred p;\n"
        "mov.b64 da, {%1, %2};\n"
        "mov.b64 db, {%3, %4};\n"
        "setp.ne.b32 p, %6, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n"
        "}\n" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(acc));
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
// Elected-lane forms of the producer's instructions (see "Warp-uniform issue" below: UTMALDG takes its tensor-map address
// and coordinates from uniform registers too).
__device__ __forceinline__ void mbar_expect_tx_e(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n@e mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n}\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_e(uint32_t bar) {
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n@e mbarrier.arrive.shared::cta.b64 _, [%0];\n}\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tma_load_4d_e(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2, int c3) {
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n"
                 "@e cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];\n}\n"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_load_3d_e(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile("{\n.reg .pred e;\nelect.sync _|e, 0xffffffff;\n"
                 "@e cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];\n}\n"
                 ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
// Warp-uniform issue: ALL lanes of the issuing warp run the loop (barrier waits, descriptor arithmetic) with identical values
// and one elected lane issues the instruction.  In a block entered by `if (lane == 0)` the compiler cannot know that the
// descriptors are warp-uniform and moves every one of them to the uniform registers UTCHMMA wants through an
// ELECT / R2UR.BROADCAST / BRA.U.ANY loop - four such moves per MMA.
template <int ACC>
__device__ __forceinline__ void umma_bf16_lh_e(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc) {
    asm volatile(
        "{\n"
        ".reg .b64 da, db;\n"
        ".reg .pred p, e;\n"
        "mov.b64 da, {%1, %2};\n"
        "mov.b64 db, {%3, %4};\n"
        "setp.ne.b32 p, %6, 0;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n"
        "}\n" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "n"(ACC));
}
__device__ __forceinline__ void umma_bf16_lhr_e(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n"
        ".reg .b64 da, db;\n"
        ".reg .pred p, e;\n"
        "mov.b64 da, {%1, %2};\n"
        "mov.b64 db, {%3, %4};\n"
        "setp.ne.b32 p, %6, 0;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n"
        "}\n" ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(acc));
}
__device__ __forceinline__ void umma_commit_e(uint32_t bar) {
    asm volatile(
        "{\n"
        ".reg .pred e;\n"
        "elect.sync _|e, 0xffffffff;\n"
        "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n"
        "}\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand descriptor for a tile of rows x (64|32) bf16, rows `row_bytes` apart, 8-row swizzle
// atoms packed densely (SBO = 8 * row_bytes).  cute::UMMA::SmemDescriptor bit layout.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t row_bytes) {
    const uint64_t layout = row_bytes == 128 ? 2ull : 4ull;       // SWIZZLE_128B : SWIZZLE_64B
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)((8 * row_bytes) >> 4) << 32) | (1ull << 46) | (layout << 61);
}

__global__ void __launch_bounds__(kHaloThreads, 1)
conv_tc_kernel(const __grid_constant__ PhaseMaps tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ PhaseMaps tmA_tail, const __grid_constant__ CUtensorMap tmB_tail,
               bf16* __restrict__ dst, const ConvEpilogue ep, const __grid_constant__ TcParams p) {
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;             // swizzle atoms need 1024 B alignment
    const uint32_t sA = base, sB = base + p.stages * p.a_stage;
    const uint32_t bars = sB + p.stages * p.b_stage;                         // 8 B each
    const uint32_t bar_full = bars, bar_empty = bars + 8 * kMaxStages;
    const uint32_t bar_tfull = bars + 16 * kMaxStages, bar_tempty = bar_tfull + 16;
    const uint32_t tmem_slot = bar_tempty + 16;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // fused BatchNorm statistics: [8 epilogue warps][sum, sum of squares][nca columns] running doubles, then one flag
    const int nca = (p.ncp + 31) & ~31;
    double* sstat = reinterpret_cast<double*>(smem_raw + (tmem_slot + 16 - smem_u32(smem_raw)));
    int* sflag = reinterpret_cast<int*>(sstat + 16 * nca);
    if (ep.stats)
        for (int i = threadIdx.x; i < 16 * nca; i += blockDim.x) sstat[i] = 0.0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(bar_tfull + 8 * a, 1); mbar_init(bar_tempty + 8 * a, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)p.tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_wait();               // everything above touched only shared memory, TMEM and kernel parameters (common.cuh)
    pdl_trigger();

    const int taps = p.taps.n;
    const int tiles_per_img = p.tiles_x * p.tiles_y;
    const int m_tiles = tiles_per_img * p.N;

    if (warp == 0) {
        // ================= TMA producer =================
        {   // all 32 lanes run the loop, one elected lane issues (*_e)
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < p.total_tiles; t += gridDim.x) {
                const int nt = t / m_tiles, mt = t - nt * m_tiles;
                const int img = mt / tiles_per_img, rem = mt - img * tiles_per_img;
                const int ty = rem / p.tiles_x;
                const int y0 = ty * p.th, x0 = (rem - ty * p.tiles_x) * p.tw;
                for (int tap = 0; tap < taps; ++tap) {
                    const int sy = y0 + p.taps.dy[tap], sx = x0 + p.taps.dx[tap];
                    const CUtensorMap* mapA = &tmA.m[p.taps.map[tap]];
                    const CUtensorMap* mapAt = &tmA_tail.m[p.taps.map[tap]];
                    const int wtap = p.taps.wtap[tap];
                    for (int kc = 0; kc < p.kchunks; ++kc) {
                        mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                        if (p.tail && kc == p.kchunks - 1) {
                            // short channel tail (72 = 64 + 8, 136 = 128 + 8, ...): a 16- or 32-channel box with the
                            // matching narrower swizzle instead of a zero-filled 64-channel one - a quarter / half of the fill
                            mbar_expect_tx_e(bar_full + 8 * stage, p.a_bytes_tail + p.b_bytes_tail);
                            tma_load_4d_e(sA + stage * p.a_stage, mapAt, bar_full + 8 * stage, kc * p.chunk, sx, sy, img);
                            tma_load_3d_e(sB + stage * p.b_stage, &tmB_tail, bar_full + 8 * stage, kc * p.chunk, nt * p.BN, wtap);
                        } else {
                            mbar_expect_tx_e(bar_full + 8 * stage, p.a_bytes + p.b_bytes);
                            tma_load_4d_e(sA + stage * p.a_stage, mapA, bar_full + 8 * stage, kc * p.chunk, sx, sy, img);
                            tma_load_3d_e(sB + stage * p.b_stage, &tmB, bar_full + 8 * stage, kc * p.chunk, nt * p.BN, wtap);
                        }
                        if (++stage == p.stages) { stage = 0; phase ^= 1; }
                    }
                }
            }
            pdl_trigger_late();
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        {   // all 32 lanes run the loop, one elected lane issues (umma_*_e)
            // cute::UMMA::InstrDescriptor: c=F32 (bit 4), a=b=BF16 (bits 7, 10), K-major A and B, N>>3 @17, M>>4 @24
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((128u >> 4) << 24);
            const uint32_t row_bytes = p.chunk * 2;
            const uint32_t hi = (uint32_t)((8 * row_bytes) >> 4) | (1u << 14) | ((row_bytes == 128 ? 2u : 4u) << 29);
            const uint32_t a_lo0 = ((sA & 0x3FFFFu) >> 4) | (1u << 16), b_lo0 = ((sB & 0x3FFFFu) >> 4) | (1u << 16);
            const uint32_t a_step = p.a_stage >> 4, b_step = p.b_stage >> 4;
            const int full_ksteps = p.chunk / 16;
            const int last_ksteps = (p.Kp - (p.kchunks - 1) * p.chunk + 15) / 16;
            // tail chunk: rows of 32 B (SWIZZLE_32B, layout 6) or 64 B (SWIZZLE_64B, layout 4), 8-row groups densely packed
            const uint32_t tail_row = (uint32_t)p.tail * 2;
            const uint32_t hi_tail = (uint32_t)((8 * tail_row) >> 4) | (1u << 14) | ((tail_row == 32 ? 6u : 4u) << 29);
            const int kchunks = p.kchunks, stages = p.stages, BN = p.BN, total = p.total_tiles;
            int stage = 0;
            uint32_t phase = 0;
            int it = 0;
            for (int t = blockIdx.x; t < total; t += gridDim.x, ++it) {
                const int as = it & 1;
                mbar_wait(bar_tempty + 8 * as, ((it >> 1) & 1) ^ 1);         // epilogue has drained this accumulator
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + as * BN;
                uint32_t acc = 0;
                for (int tap = 0; tap < taps; ++tap) {
                    for (int kc = 0; kc < kchunks; ++kc) {
                        mbar_wait(bar_full + 8 * stage, phase);              // TMA bytes have landed
                        tc_fence_after();
                        const bool is_tail = p.tail && kc == kchunks - 1;
                        const int ksteps = kc == kchunks - 1 ? last_ksteps : full_ksteps;
                        const uint32_t hik = is_tail ? hi_tail : hi;
                        const uint32_t da = a_lo0 + stage * a_step, db = b_lo0 + stage * b_step;
#pragma unroll
                        for (int k = 0; k < 4; ++k)
                            if (k < ksteps) { umma_bf16_lhr_e(tmem_d, da + 2 * k, hik, db + 2 * k, hik, idesc, acc); acc = 1; }   // +32 B per K=16 step
                        umma_commit_e(bar_empty + 8 * stage);                  // frees the smem stage when the MMAs retire
                        if (++stage == stages) { stage = 0; phase ^= 1; }
                    }
                }
                umma_commit_e(bar_tfull + 8 * as);                             // accumulator complete -> epilogue
            }
        }
    } else {
        // ================= epilogue: two groups of four warps, group g drains accumulator g =================
        const int quad = warp & 3, grp = (warp - 2) >> 2;                    // TMEM lane quadrant this warp may read
        const int m = quad * 32 + lane;
        const int dy = m / p.tw, dx = m % p.tw;
        const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + grp * p.BN;
        int it = grp;
        for (int t = blockIdx.x + grp * gridDim.x; t < p.total_tiles; t += 2 * gridDim.x, it += 2) {
            const int nt = t / m_tiles, mt = t - nt * m_tiles;
            const int img = mt / tiles_per_img, rem = mt - img * tiles_per_img;
            const int ty = rem / p.tiles_x;
            const int y = ty * p.th + dy, x = (rem - ty * p.tiles_x) * p.tw + dx;
            const bool valid = dy < p.th && y < p.H && x < p.W;
            const int n0 = nt * p.BN;
            const int ncols = min(p.BN, p.ncp - n0);                         // multiple of 8
            bf16* out = dst + ((int64_t)(img * p.Hd + y * p.dmul + p.da) * p.Wd + x * p.dmul + p.db) * p.ldd + n0;
            mbar_wait(bar_tfull + 8 * grp, (it >> 1) & 1);
            tc_fence_after();
            double* ws_s = sstat + ((warp - 2) * 2 + 0) * nca + n0;       // rows cover every output column, whichever n tiles this CTA visits
            double* ws_q = sstat + ((warp - 2) * 2 + 1) * nca + n0;
            int c0 = 0;
            for (; c0 < ncols && ncols - c0 > 16; c0 += 32) {
                uint32_t v[32];
                tmem_ld32(taddr + c0, v);
                tmem_ld_wait();
                epilogue_chunk<32>(v, n0 + c0, ncols - c0, valid, out + c0, p.Nc, ep, ws_s + c0, ws_q + c0, lane);
            }
            for (; c0 < ncols; c0 += 16) {
                uint32_t v[16];
                tmem_ld16(taddr + c0, v);
                tmem_ld_wait();
                epilogue_chunk<16>(v, n0 + c0, ncols - c0, valid, out + c0, p.Nc, ep, ws_s + c0, ws_q + c0, lane);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + 8 * grp);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)p.tmem_cols));
    }
    if (ep.stats) stats_tail(ep, sstat, nca, sflag);
}

// =================================================================================================
// "Halo" variant for layers whose whole reduction fits one 64-channel chunk and whose weights fit in
// shared memory (DUCK-Net's 17/34-channel levels, UNet's 32/64: >80 % of the conv bytes).
// The streaming kernel above re-fetches the A tile once per tap (9x L2->SM traffic for a 3x3).  Here the
// tile is an 8 (x) by 16 (y) pixel patch and ONE TMA box brings the patch plus its halo, Wh = 8 + tap reach
// pixels wide, into shared memory.  The 128B swizzle is a function of the absolute shared-memory address
// (measured), so the A operand of tap (r, s) is the same shared memory seen through a descriptor whose start
// is shifted by (r'*Wh + s') pixel rows and whose 8-row groups are Wh*128 B apart - whatever Wh is: the 8
// pixels of tile row dy are 8 consecutive halo rows.  All taps' weights stay resident.
// Measured cost model of this kernel at 105 tiles per CTA (17->17 and 32->32 layers at 352x352):
//   one MMA issuer:  t/tile ~ 600 cycles of barrier round trips + ~57 cycles per 128xNx16 MMA with N = 32
//                    (SS-mode operand fetch: 4 KB of A from shared memory per MMA, not the 16-cycle tensor floor)
//   two issuers on alternate tiles hide the round trips (17->17 3x3: 113 -> 85 us, 32->32: 109 -> 75 us).  What is left
//   tracks the number of box rows (~4 cycles per pixel row of the halo box) and the shared-memory traffic of the MMAs;
//   a second TMA-issuing thread and two-tile boxes (pair mode) did not move it.
struct HaloParams {
    int N, H, W, tiles_x, tiles_y, total_tiles;
    int kh, kw, mul_t, off_h, off_w, oy_min, ox_min;
    int Wh, Hh, BN, ksteps, Nc, ncp, ldd;
    int stages, tmem_cols, nacc;   // nacc TMEM accumulators in flight (hides the commit -> epilogue -> release round trip)
    uint32_t a_bytes, a_stage, b_tap_bytes;
    // channel tail (68 channels = 64 + a 16-wide box, SWIZZLE_32B): a second, 32 B-per-pixel patch after the main one in
    // every stage and a second set of resident weights after the main ones; one more K step per tap
    uint32_t a_tail_off, a_bytes_tail, b_tail_bytes, b_tail_base;
    uint32_t rb;                   // bytes of one pixel row in shared memory: 128 (64-channel box, SWIZZLE_128B) or, for K <= 32
                                   // channels, 64 (32-channel box, SWIZZLE_64B): the TMA unit's time per row follows the bytes it writes
    int flat;                      // 1x1, no padding: the activation is walked as one row of N*H*W pixels, a tile = 128 consecutive
                                   // pixels = ONE contiguous run for the TMA unit instead of 16 runs of 8 pixels (tma_rate.cu: the unit
                                   // pays ~20-35 cycles per run on top of ~1.2-1.5 per row)
    int issuers;                   // 1 or 2 MMA-issuing warps
    int tw, boxes_x, total_boxes;  // tw == 2 (pair mode): one TMA box covers two tiles side by side; issuer / epilogue group j takes tile j
#ifdef MSX_ABLATION
    int dbg;                       // MSX_DBG_HALO ablations (results are wrong): 1 no stores, 2 no MMAs, 4 no TMA loads, 8 no tcgen05.ld
#endif
};
// Ablation switches exist only in -DMSX_ABLATION builds (tools/microbench.py measurements); the shipped kernel has none.
#ifdef MSX_ABLATION
#define MSX_DBG(p, bit) ((p).dbg & (bit))
#else
#define MSX_DBG(p, bit) 0
#endif

// Start address shifted by whole 128 B pixel rows, base_offset field left 0: measured on B200, the
// 128B swizzle is applied to the absolute shared-memory address bits, so a window that starts at any
// row of a TMA-written (1024 B aligned) buffer reads back consistently.  (Setting the descriptor's
// base_offset to (addr >> 7) & 7 gives wrong results.)
__device__ __forceinline__ uint64_t umma_desc_halo(uint32_t saddr, uint32_t sbo_bytes) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | (1ull << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

template <int TAPS, int KSTEPS, bool TAIL>
__global__ void __launch_bounds__(kHaloThreads2, 1)
conv_tc_halo_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmAt, const __grid_constant__ CUtensorMap tmBt,
                    bf16* __restrict__ dst, const ConvEpilogue ep, const HaloParams p) {
    extern __shared__ uint8_t smem_raw[];
    const int taps = p.kh * p.kw;
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t sB = base, sBt = base + p.b_tail_base;
    const uint32_t sA = base + ((p.b_tail_base + (TAIL ? taps * p.b_tail_bytes : 0u) + 1023u) & ~1023u);
    const uint32_t bars = sA + p.stages * p.a_stage;
    const uint32_t bar_full = bars, bar_empty = bars + 8 * kMaxStages;
    const uint32_t bar_tfull = bars + 16 * kMaxStages, bar_tempty = bar_tfull + 8 * kMaxAcc, bar_w = bar_tempty + 8 * kMaxAcc;
    const uint32_t tmem_slot = bar_w + 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // fused BatchNorm statistics (see conv_tc_kernel)
    const int nca = (p.ncp + 31) & ~31;
    double* sstat = reinterpret_cast<double*>(smem_raw + (tmem_slot + 24 - smem_u32(smem_raw)));
    int* sflag = reinterpret_cast<int*>(sstat + 16 * nca);
    if (ep.stats)
        for (int i = threadIdx.x; i < 16 * nca; i += blockDim.x) sstat[i] = 0.0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, p.tw == 2 ? 2 : 1); }   // pair mode: both issuers release a box
        for (int a = 0; a < p.nacc; ++a) { mbar_init(bar_tfull + 8 * a, 1); mbar_init(bar_tempty + 8 * a, 4); }
        mbar_init(bar_w, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)p.tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_wait();               // everything above touched only shared memory, TMEM and kernel parameters (common.cuh)
    pdl_trigger();
    // work unit of the producer: one tile (8 x 16 pixels) or, in pair mode, a box of two tiles side by side (16 x 16)
    const bool pair = p.tw == 2;
    const int units_x = pair ? p.boxes_x : p.tiles_x, unit_w = pair ? 16 : 8;
    const int tiles_per_img = units_x * p.tiles_y, total_units = pair ? p.total_boxes : p.total_tiles;

    if (warp == 0) {
        // (a second TMA-issuing warp on alternate tiles was measured: no gain - the per-box cost is in the TMA unit)
        {   // all 32 lanes run the loop, one elected lane issues (*_e)
            mbar_expect_tx_e(bar_w, taps * (p.b_tap_bytes + (TAIL ? p.b_tail_bytes : 0u)));      // resident weights, once per CTA
            for (int tap = 0; tap < taps; ++tap) tma_load_3d_e(sB + tap * p.b_tap_bytes, &tmB, bar_w, 0, 0, tap);
            if (TAIL)
                for (int tap = 0; tap < taps; ++tap) tma_load_3d_e(sBt + tap * p.b_tail_bytes, &tmBt, bar_w, 64, 0, tap);
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < total_units; t += gridDim.x) {
                const int img = t / tiles_per_img, rem = t - img * tiles_per_img;
                const int ty = rem / units_x;
                const int y0 = ty * 16, x0 = p.flat ? rem * 128 : (rem - ty * units_x) * unit_w;
                mbar_wait(bar_empty + 8 * stage, phase ^ 1);
                if (MSX_DBG(p, 4)) mbar_arrive_e(bar_full + 8 * stage);
                else {
                    mbar_expect_tx_e(bar_full + 8 * stage, p.a_bytes + (TAIL ? p.a_bytes_tail : 0u));
                    tma_load_4d_e(sA + stage * p.a_stage, &tmA, bar_full + 8 * stage, 0, x0 + p.ox_min, y0 + p.oy_min, img);
                    if (TAIL) tma_load_4d_e(sA + stage * p.a_stage + p.a_tail_off, &tmAt, bar_full + 8 * stage, 64, x0 + p.ox_min, y0 + p.oy_min, img);
                }
                if (++stage == p.stages) { stage = 0; phase ^= 1; }
            }
            pdl_trigger_late();
        }
    } else if (warp == 1 || warp == 10) {
        // Two issuers (p.issuers == 2) take alternate tiles: the per-tile barrier round trips of one overlap the
        // tensor-pipe back-pressure of the other.  Accumulators and stages are per tile, so no ordering between them.
        const int me = warp == 1 ? 0 : 1;
        if (me < p.issuers) {   // all 32 lanes run the loop, one elected lane issues (umma_*_e)
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.BN >> 3) << 17) | ((128u >> 4) << 24);
            // The issuing thread is latency bound on its own instruction stream (measured: the issue loop, not
            // TMA, MMA or the epilogue, set the tile rate), so TAPS / KSTEPS are template constants, the loops
            // are fully unrolled without predicates and per-tap address offsets live in registers.
            uint32_t a_off[TAPS], b_lo[TAPS], a_off_t[TAIL ? TAPS : 1], b_lo_t[TAIL ? TAPS : 1];
#pragma unroll
            for (int tap = 0; tap < TAPS; ++tap) {
                const int r = tap / p.kw, s = tap - r * p.kw;
                const int ry = r * p.mul_t + p.off_h - p.oy_min, sx = s * p.mul_t + p.off_w - p.ox_min;
                a_off[tap] = (uint32_t)((ry * p.Wh + sx) * (int)p.rb) >> 4;
                b_lo[tap] = (((sB + tap * p.b_tap_bytes) & 0x3FFFFu) >> 4) | (1u << 16);
                if (TAIL) {
                    a_off_t[tap] = (uint32_t)((ry * p.Wh + sx) * 32) >> 4;          // 32 B per pixel row in the tail patch
                    b_lo_t[tap] = (((sBt + tap * p.b_tail_bytes) & 0x3FFFFu) >> 4) | (1u << 16);
                }
            }
            const uint32_t hi_a_t = (uint32_t)((p.Wh * 32) >> 4) | (1u << 14) | (6u << 29);     // SBO, version 1, SWIZZLE_32B
            const uint32_t hi_b_t = (uint32_t)(256 >> 4) | (1u << 14) | (6u << 29);
            const uint32_t lay = p.rb == 128 ? 2u : 4u;                                        // SWIZZLE_128B : SWIZZLE_64B
            const uint32_t hi_a = (uint32_t)((p.Wh * p.rb) >> 4) | (1u << 14) | (lay << 29);     // SBO, version 1, swizzle
            const uint32_t hi_b = (uint32_t)((8 * p.rb) >> 4) | (1u << 14) | (lay << 29);
            const int nacc = p.nacc, stages = p.stages, total = total_units, BN = p.BN;
            // pair mode: both issuers walk every box; issuer `me` computes the tile at x offset 8 * me (8 pixel rows)
            const uint32_t a_stage16 = p.a_stage >> 4, a_lo0 = (((sA & 0x3FFFFu) >> 4) | (1u << 16)) + (pair ? me * (int)((8 * p.rb) >> 4) : 0);
            mbar_wait(bar_w, 0);
            tc_fence_after();
            const int step = pair ? 1 : p.issuers, astep = pair ? 2 : p.issuers;
            int stage = pair ? 0 : me % stages, as = me % nacc;
            uint32_t phase = pair ? 0 : (me / stages) & 1, aphase = 1 ^ ((me / nacc) & 1);
            for (int t = pair ? blockIdx.x : blockIdx.x + me * gridDim.x; t < total; t += step * gridDim.x) {
                mbar_wait(bar_tempty + 8 * as, aphase);
                mbar_wait(bar_full + 8 * stage, phase);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + as * BN;
                const uint32_t a_lo = a_lo0 + stage * a_stage16;
                if (!MSX_DBG(p, 2)) {
#pragma unroll
                for (int tap = 0; tap < TAPS; ++tap) {
#pragma unroll
                    for (int k = 0; k < KSTEPS; ++k) {
                        if (tap == 0 && k == 0) umma_bf16_lh_e<0>(tmem_d, a_lo + a_off[tap], hi_a, b_lo[tap], hi_b, idesc);
                        else umma_bf16_lh_e<1>(tmem_d, a_lo + a_off[tap] + 2 * k, hi_a, b_lo[tap] + 2 * k, hi_b, idesc);
                    }
                    if (TAIL) umma_bf16_lh_e<1>(tmem_d, a_lo + (p.a_tail_off >> 4) + a_off_t[tap], hi_a_t, b_lo_t[tap], hi_b_t, idesc);
                }
                }
                if (MSX_DBG(p, 16)) { if (lane == 0) { mbar_arrive(bar_empty + 8 * stage); mbar_arrive(bar_tfull + 8 * as); } }
                else {
                    umma_commit_e(bar_empty + 8 * stage);
                    umma_commit_e(bar_tfull + 8 * as);
                }
                stage += step;
                if (stage >= stages) { stage -= stages; phase ^= 1; }
                as += astep;
                if (as >= nacc) { as -= nacc; aphase ^= 1; }
            }
        }
    } else {
        // ===== epilogue: two groups of four warps; group g drains accumulator g (tiles with it % 2 == g), so
        // two tiles are converted and stored concurrently (each group alone is latency bound) =====
        const int quad = warp & 3, grp = (warp - 2) >> 2;
        const int m = quad * 32 + lane;
        const int dy = m >> 3, dx = m & 7;
        const int ncols = p.ncp;
        const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
#ifdef MSX_NO_DEFER_STATS                                            // measurement build: per-tile column reduction (110 instead of 168 registers)
        const bool defer = false;
#else
        const bool defer = ep.stats && ncols <= 32;               // one chunk: statistics in registers until the end of the launch
#endif
        float ra[32], rb[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) { ra[j] = 0.f; rb[j] = 0.f; }
        int it = grp;
        for (int t = pair ? blockIdx.x : blockIdx.x + grp * gridDim.x; t < total_units; t += (pair ? 1 : 2) * gridDim.x, it += 2) {
            const int img = t / tiles_per_img, rem = t - img * tiles_per_img;
            const int ty = rem / units_x;
            const int y = p.flat ? 0 : ty * 16 + dy;
            const int x = p.flat ? rem * 128 + m : (rem - ty * units_x) * unit_w + (pair ? grp * 8 : 0) + dx;
            const bool valid = y < p.H && x < p.W && !MSX_DBG(p, 1);
            bf16* out = dst + ((int64_t)(img * p.H + y) * p.W + x) * p.ldd;
            const int as = it % p.nacc;
            const uint32_t taddr = tlane + as * p.BN;
            mbar_wait(bar_tfull + 8 * as, (it / p.nacc) & 1);
            tc_fence_after();
            double* ws_s = sstat + ((warp - 2) * 2 + 0) * nca;
            double* ws_q = sstat + ((warp - 2) * 2 + 1) * nca;
            int c0 = 0;
            if (defer) {
                uint32_t v[32];
                if (!MSX_DBG(p, 8)) {
                    if (ncols > 16) tmem_ld32(taddr, v);
                    else {
                        uint32_t v16[16];
                        tmem_ld16(taddr, v16);
#pragma unroll
                        for (int j = 0; j < 16; ++j) { v[j] = v16[j]; v[16 + j] = 0u; }
                    }
                    tmem_ld_wait();
                }
                epilogue_chunk_defer<32>(v, 0, ncols, valid, out, p.Nc, ep, ra, rb);
                c0 = ncols;
            }
            for (; c0 + 32 <= ncols || (c0 < ncols && ncols - c0 > 16); c0 += 32) {
                uint32_t v[32];
                if (!MSX_DBG(p, 8)) {
                    tmem_ld32(taddr + c0, v);
                    tmem_ld_wait();
                }
                epilogue_chunk<32>(v, c0, ncols - c0, valid, out + c0, p.Nc, ep, ws_s + c0, ws_q + c0, lane);
            }
            for (; c0 < ncols; c0 += 16) {
                uint32_t v[16];
                tmem_ld16(taddr + c0, v);
                tmem_ld_wait();
                epilogue_chunk<16>(v, c0, ncols - c0, valid, out + c0, p.Nc, ep, ws_s + c0, ws_q + c0, lane);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_tempty + 8 * as);
        }
        if (defer) stats_flush<32>(ra, rb, sstat + ((warp - 2) * 2 + 0) * nca, sstat + ((warp - 2) * 2 + 1) * nca, lane);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)p.tmem_cols));
    }
    if (ep.stats) stats_tail(ep, sstat, nca, sflag);
}

// ---- host side ----------------------------------------------------------------------------------
// cudaFuncAttributeMaxDynamicSharedMemorySize is a per-device (per-context) setting: opt in once per (device, kernel).
// Called from the caller's thread and from the autograd worker thread, hence the lock.
int ensure_big_smem(const void* kernel) {
    static std::mutex mu;
    static std::set<std::pair<int, const void*>> done;
    int dev = 0;
    MSX_CHECK_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(mu);
    if (done.count({dev, kernel})) return MSX_OK;
    MSX_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    done.insert({dev, kernel});
    return MSX_OK;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = []() -> EncodeTiledFn {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            return nullptr;
        return reinterpret_cast<EncodeTiledFn>(f);
    }();
    return fn;
}

// Encoded tensor maps are pure functions of (pointer, geometry): the eager (non-graph) path re-encounters the same few
// hundred of them every step (the caching allocator hands the same blocks back), so they are kept in a small
// direct-mapped per-thread cache instead of calling the driver's encoder twice per convolution.
struct MapKey {
    const void* ptr;
    cuuint64_t dims[4], strides[3];
    cuuint32_t box[4];
    int rank, sw;
};
struct MapSlot { MapKey key; CUtensorMap map; bool used; };
constexpr int kMapSlots = 2048;

int make_map(CUtensorMap* m, int rank, const void* ptr, const cuuint64_t* dims, const cuuint64_t* strides_bytes,
             const cuuint32_t* box, CUtensorMapSwizzle sw) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return set_error(MSX_ERR_CUDA, "cuTensorMapEncodeTiled entry point unavailable");
    static thread_local MapSlot* cache = static_cast<MapSlot*>(calloc(kMapSlots, sizeof(MapSlot)));
    MapKey key;
    memset(&key, 0, sizeof(key));
    key.ptr = ptr; key.rank = rank; key.sw = (int)sw;
    for (int i = 0; i < rank; ++i) { key.dims[i] = dims[i]; key.box[i] = box[i]; }
    for (int i = 0; i + 1 < rank; ++i) key.strides[i] = strides_bytes[i];
    uint64_t h = 1469598103934665603ull;
    for (size_t i = 0; i < sizeof(key) / 8; ++i) { h ^= reinterpret_cast<const uint64_t*>(&key)[i]; h *= 1099511628211ull; }
    MapSlot* slot = cache ? &cache[(h ^ (h >> 29)) % kMapSlots] : nullptr;
    if (slot && slot->used && memcmp(&slot->key, &key, sizeof(key)) == 0) { *m = slot->map; return MSX_OK; }
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r = fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(ptr), dims, strides_bytes, box, es,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return set_error(MSX_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    if (slot) { slot->key = key; slot->map = *m; slot->used = true; }
    return MSX_OK;
}

void pick_tile(int H, int W, int& tw, int& th) {
    double best = -1;
    tw = 1; th = 1;
    for (int w = 1; w <= (W < 128 ? W : 128); ++w) {
        int h = 128 / w;
        if (h > H) h = H;
        if (h < 1) continue;
        const double tiles = (double)((W + w - 1) / w) * ((H + h - 1) / h);
        const double eff = (double)W * H / (tiles * 128.0);
        if (eff > best + 1e-9 || (eff > best - 1e-9 && w > tw)) { best = eff; tw = w; th = h; }
    }
}


// =================================================================================================
// Weight gradient on tcgen05:  dW[a][b][tap] = sum_p small[p][a] * big[p + off(tap)][b]
//   (Conv2d stride 1: small = dy (a = cout), big = x (b = cin)).
// GEMM view: M = 128 a-channels, N = TN <= 256 b-channels, K = pixels.  Both operands are the NHWC
// tiles themselves ([128 pixel rows][64|32 channels], TMA + swizzle) used as MN-MAJOR UMMA operands:
// channels run along the 128 B (64 B) row, pixels along K, so no transpose is ever materialised.
// The box of a tile is always a full 128-pixel rectangle (tw*th == 128); whatever hangs over the image
// is TMA zero fill, so out-of-image pixels add exactly 0 to the sums.
// A CTA owns one (a tile, b tile, tap group, pixel split): it keeps tg taps x TN fp32 columns in TMEM
// (<= 512), streams its pixel tiles through two smem rings (A: dy tile, reused by all taps of the
// group; B: one shifted x tile per tap) and finally dumps TMEM to the split-K workspace that
// wgrad_reduce_kernel folds into the reference OIHW layout.
struct WgParams {
    int N, H, W, tw, th, tiles_x, tiles_y;
    int kh, kw, dil, pad_h, pad_w;
    int a_tiles, b_tiles, TN, tg, tap_groups, splits;
    int chunkA, chunkB, groupsB;      // channels per MN group, groups per B tile
    int Ap, Bp;                       // round_up(A, 8), round_up(B, 8): tensor-map channel extents
    int Apad, Bpad;                   // workspace leading dims (a_tiles*128, b_tiles*TN)
    int stagesB, tmem_cols, SA;
    uint32_t a_stage, b_stage;
    int64_t pix_tiles;
    int halo, Wh, Hh;                 // halo mode: one x patch (+halo) per pixel tile serves every tap
    TapTable taps;                    // streaming mode: per tap, which phase view of the big side and which shift
    int TNa, G, ncg;                  // tap-stacked mode (halo == 2): N = dy channels, G x-taps share one MMA's M axis
    int gs;                           // pixels between the G stacked windows: dil, or 0 for a one-column filter (kw == 1: the groups
                                      // past the first repeat it and are never read back, and the patch needs no x halo)
    // halo == 2 with rowN > 0: the kh filter ROWS are stacked on the MMA's N axis as well.  The x patch then carries no
    // y halo; the dy patch does (HhA = 16 + (kh-1)*dil rows) and is the MN-major B operand whose N groups - rowN channels
    // each - are dil image rows apart (block j <-> filter row kh-1-j).  One MMA covers G x kh taps: a 3x3 at <= 32
    // channels needs 8 MMAs per 128-pixel tile instead of 24.
    int rowN, HhA;
    // halo == 2 with wide > 0: 16 x 8-pixel tiles instead of 8 x 16.  The TMA unit pays per contiguous run of a box (~20-35
    // cycles, tools/micro/tma_rate.cu) and a box has one run per image row: half as many rows, twice as long.  For the MN-major
    // operands only the K order changes - one K = 16 step is one image row of the tile (two dense 8-pixel groups) instead of two.
    int wide;
    uint32_t a_box_bytes;             // bytes of one dy TMA box
#ifdef MSX_ABLATION
    int dbg;                          // MSX_DBG_WGRAD ablations of the stacked (halo == 2) mode, results are wrong: 2 no MMAs, 4 no TMA loads
#endif
};

// MN-major operand descriptor: MN groups `lbo` bytes apart, 8-row (K) groups 8*row_bytes apart.
__device__ __forceinline__ uint64_t umma_desc_mn(uint32_t saddr, uint32_t row_bytes, uint32_t lbo) {
    const uint64_t layout = row_bytes == 128 ? 2ull : 4ull;
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
           ((uint64_t)((8 * row_bytes) >> 4) << 32) | (1ull << 46) | (layout << 61);
}

__global__ void __launch_bounds__(kThreads, 1)
wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmS, const __grid_constant__ PhaseMaps tmG,
                float* __restrict__ partial, const __grid_constant__ WgParams p) {
    extern __shared__ uint8_t smem_raw[];
    const int SA = p.SA;                                   // dy-tile ring depth (2, or 4 when the tiles are small)
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t sA = base, sB = base + SA * p.a_stage;
    const uint32_t bars = sB + p.stagesB * p.b_stage;
    const uint32_t barA_full = bars, barA_empty = bars + 32, barB_full = bars + 64, barB_empty = barB_full + 8 * kMaxStages;
    const uint32_t bar_done = barB_empty + 8 * kMaxStages, tmem_slot = bar_done + 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // unit decode: blockIdx.x = ((split * tap_groups + tgi) * b_tiles + bt) * a_tiles + at
    int u = blockIdx.x;
    const int at = u % p.a_tiles; u /= p.a_tiles;
    const int bt = u % p.b_tiles; u /= p.b_tiles;
    const int tgi = u % p.tap_groups; u /= p.tap_groups;
    const int split = u;
    const int taps = p.kh * p.kw;
    const int tap0 = tgi * p.tg, ntap = min(p.tg, taps - tap0);
    const int64_t per = (p.pix_tiles + p.splits - 1) / p.splits;
    const int64_t t_beg = (int64_t)split * per, t_end = min(t_beg + per, p.pix_tiles);
    const int a0 = at * 128, b0 = bt * p.TN;
    const int groupsA = min((min(p.Ap - a0, 128) + p.chunkA - 1) / p.chunkA, 128 / p.chunkA);
    const uint32_t rbA = p.chunkA * 2, rbB = p.chunkB * 2;

    if (threadIdx.x == 0) {
        for (int s = 0; s < SA; ++s) { mbar_init(barA_full + 8 * s, 1); mbar_init(barA_empty + 8 * s, 1); }
        for (int s = 0; s < p.stagesB; ++s) { mbar_init(barB_full + 8 * s, 1); mbar_init(barB_empty + 8 * s, 1); }
        mbar_init(bar_done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"((uint32_t)p.tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    pdl_wait();               // everything above touched only shared memory, TMEM and kernel parameters (common.cuh)
    pdl_trigger();
    const int64_t tiles_per_img = (int64_t)p.tiles_x * p.tiles_y;

    if (warp == 0) {
        {   // all 32 lanes run the loop, one elected lane issues (*_e)
            int stA = 0, stB = 0;
            uint32_t phA = 0, phB = 0;
            // tile coordinates by increments: the 64-bit divisions of t cost more than the rest of an iteration
            int img_i = (int)(t_beg / tiles_per_img), tyi, txi;
            {
                const int rem0 = (int)(t_beg % tiles_per_img);
                tyi = rem0 / p.tiles_x; txi = rem0 - tyi * p.tiles_x;
            }
            for (int64_t t = t_beg; t < t_end; ++t) {
                const int img = img_i, y0 = tyi * p.th, x0 = txi * p.tw;
                if (++txi == p.tiles_x) { txi = 0; if (++tyi == p.tiles_y) { tyi = 0; ++img_i; } }
                const int ya = p.rowN ? y0 - (p.kh - 1) * p.dil + p.pad_h : y0;        // rows stacked on N: dy patch with the y halo
                mbar_wait(barA_empty + 8 * stA, phA ^ 1);
                if (MSX_DBG(p, 4)) mbar_arrive_e(barA_full + 8 * stA);
                else {
                    mbar_expect_tx_e(barA_full + 8 * stA, groupsA * p.a_box_bytes);
                    for (int g = 0; g < groupsA; ++g)
                        tma_load_4d_e(sA + stA * p.a_stage + g * 128 * rbA, &tmS, barA_full + 8 * stA, a0 + g * p.chunkA, x0, ya, img);
                }
                if (++stA == SA) { stA = 0; phA ^= 1; }
                if (p.halo) {
                    mbar_wait(barB_empty + 8 * stB, phB ^ 1);
                    if (MSX_DBG(p, 4)) mbar_arrive_e(barB_full + 8 * stB);
                    else {
                        mbar_expect_tx_e(barB_full + 8 * stB, (uint32_t)p.Hh * p.Wh * rbB);
                        tma_load_4d_e(sB + stB * p.b_stage, &tmG.m[0], barB_full + 8 * stB, b0, x0 - p.pad_w, p.rowN ? y0 : y0 - p.pad_h, img);
                    }
                    if (++stB == p.stagesB) { stB = 0; phB ^= 1; }
                } else {
                    for (int k = 0; k < ntap; ++k) {
                        const int tap = tap0 + k;
                        const CUtensorMap* mapG = &tmG.m[p.taps.map[tap]];
                        mbar_wait(barB_empty + 8 * stB, phB ^ 1);
                        mbar_expect_tx_e(barB_full + 8 * stB, p.groupsB * 128 * rbB);
                        for (int g = 0; g < p.groupsB; ++g)
                            tma_load_4d_e(sB + stB * p.b_stage + g * 128 * rbB, mapG, barB_full + 8 * stB, b0 + g * p.chunkB,
                                        x0 + p.taps.dx[tap], y0 + p.taps.dy[tap], img);
                        if (++stB == p.stagesB) { stB = 0; phB ^= 1; }
                    }
                }
            }
            pdl_trigger_late();
        }
    } else if (warp == 1) {
        {   // all 32 lanes, one elected lane issues (umma_*_e)
            // c=F32, a=b=BF16, A and B MN-major (bits 15, 16), N>>3 @17, M>>4 @24
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) |
                                   ((uint32_t)(p.TN >> 3) << 17) | ((128u >> 4) << 24);
            int stA = 0, stB = 0;
            uint32_t phA = 0, phB = 0;
            if (p.halo == 2) {
                // Tap-stacked wgrad for few-channel layers.  The x patch (+halo) is the A operand, MN-major with its
                // MN groups only ONE PIXEL apart (LBO = dil * row bytes): the G groups of one MMA are the same window
                // shifted by 0..G-1 taps in x, so M = 128 carries G taps x (32|64) channels instead of 17 useful rows,
                // and a 3x3 needs 3 MMAs per K-step instead of 9.  D[(s, ci)][co] lives in kh * ncg accumulators.
                const int accN = p.rowN ? p.kh * p.rowN : p.TNa;                   // columns of one accumulator
                const uint32_t idesc2 = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) |
                                        ((uint32_t)(accN >> 3) << 17) | ((128u >> 4) << 24);
                constexpr int kMaxAccQ = 9;
                const int nq = p.rowN ? p.ncg : p.kh * p.ncg;
                uint32_t x_off[kMaxAccQ];
#pragma unroll
                for (int q = 0; q < kMaxAccQ; ++q) {
                    const int r = p.rowN ? 0 : q / p.ncg, cg = p.rowN ? q : q - r * p.ncg;
                    x_off[q] = (uint32_t)((r * p.dil * p.Wh + cg * p.G * p.gs) * (int)rbB) >> 4;
                }
                const uint32_t hi_x = (uint32_t)(((p.wide ? 8 : p.Wh) * rbB) >> 4) | (1u << 14) | ((rbB == 128 ? 2u : 4u) << 29);
                const uint32_t hi_y = ((8 * rbA) >> 4) | (1u << 14) | ((rbA == 128 ? 2u : 4u) << 29);
                const uint32_t x_lo0 = ((sB & 0x3FFFFu) >> 4) | ((((uint32_t)p.gs * rbB) >> 4) << 16);
                // dy operand: N groups 128 pixel rows apart (several channel groups of one tile), or - rows stacked on N -
                // dil image rows (8 pixels each) apart: the same channels seen through kh shifted windows
                const uint32_t y_lbo = p.rowN ? (uint32_t)(p.dil * p.tw) * rbA : 128u * rbA;
                const uint32_t y_lo0 = ((sA & 0x3FFFFu) >> 4) | (((y_lbo >> 4) & 0x3FFFu) << 16);
                const uint32_t x_kk = (uint32_t)((p.wide ? 1 : 2) * p.Wh * rbB) >> 4, y_kk = (16 * rbA) >> 4;
                const uint32_t a_st = p.a_stage >> 4, b_st = p.b_stage >> 4;
                const int stagesB = p.stagesB, TNa = accN;
                for (int64_t t = t_beg; t < t_end; ++t) {
                    mbar_wait(barA_full + 8 * stA, phA);
                    mbar_wait(barB_full + 8 * stB, phB);
                    tc_fence_after();
                    const uint32_t y_lo = y_lo0 + stA * a_st;
                    const uint32_t x_lo = x_lo0 + stB * b_st;
                    const uint32_t first = t > t_beg ? 1u : 0u;
                    if (MSX_DBG(p, 2)) {
                        mbar_arrive_e(barB_empty + 8 * stB);
                        mbar_arrive_e(barA_empty + 8 * stA);
                    } else {
#pragma unroll
                    for (int q = 0; q < kMaxAccQ; ++q) {
                        if (q < nq) {
                            const uint32_t tmem_d = tmem_base + q * TNa;
                            umma_bf16_lhr_e(tmem_d, x_lo + x_off[q], hi_x, y_lo, hi_y, idesc2, first);
#pragma unroll
                            for (int kk = 1; kk < 8; ++kk)
                                umma_bf16_lh_e<1>(tmem_d, x_lo + x_off[q] + kk * x_kk, hi_x, y_lo + kk * y_kk, hi_y, idesc2);
                        }
                    }
                    umma_commit_e(barB_empty + 8 * stB);
                    umma_commit_e(barA_empty + 8 * stA);
                    }
                    if (++stB == stagesB) { stB = 0; phB ^= 1; }
                    if (++stA == SA) { stA = 0; phA ^= 1; }
                }
            } else if (p.halo) {
                // per-tap start offsets of the shifted x window inside the halo patch (16 B units), hoisted
                constexpr int kMaxTaps = 9;
                uint32_t b_off[kMaxTaps];
#pragma unroll
                for (int k = 0; k < kMaxTaps; ++k) {
                    const int tap = tap0 + k, r = tap / p.kw, s = tap - r * p.kw;
                    b_off[k] = (uint32_t)((r * p.dil * p.Wh + s * p.dil) * 128) >> 4;
                }
                // descriptor halves: low = start address | LBO << 16, high = SBO | version | layout
                const uint32_t lbo_a = (((128 * rbA) >> 4) & 0x3FFFu) << 16;
                const uint32_t hi_a = ((8 * rbA) >> 4) | (1u << 14) | ((rbA == 128 ? 2u : 4u) << 29);
                const uint32_t hi_b = (uint32_t)((p.Wh * 128) >> 4) | (1u << 14) | (2u << 29);                 // one MN group: LBO unused
                const uint32_t a_kk = (16 * rbA) >> 4, b_kk = (uint32_t)(2 * p.Wh * 128) >> 4;               // 16 pixels = 2 tile rows
                const uint32_t a_lo0 = ((sA & 0x3FFFFu) >> 4) | lbo_a, b_lo0 = (sB & 0x3FFFFu) >> 4;
                const uint32_t a_st = p.a_stage >> 4, b_st = p.b_stage >> 4;
                const int stagesB = p.stagesB, TN = p.TN;
                for (int64_t t = t_beg; t < t_end; ++t) {
                    mbar_wait(barA_full + 8 * stA, phA);
                    mbar_wait(barB_full + 8 * stB, phB);
                    tc_fence_after();
                    const uint32_t a_lo = a_lo0 + stA * a_st;
                    const uint32_t b_lo = b_lo0 + stB * b_st;
                    const uint32_t first = t > t_beg ? 1u : 0u;
#pragma unroll
                    for (int k = 0; k < kMaxTaps; ++k) {
                        if (k < ntap) {
                            const uint32_t tmem_d = tmem_base + k * TN;
                            umma_bf16_lhr_e(tmem_d, a_lo, hi_a, b_lo + b_off[k], hi_b, idesc, first);
#pragma unroll
                            for (int kk = 1; kk < 8; ++kk)
                                umma_bf16_lh_e<1>(tmem_d, a_lo + kk * a_kk, hi_a, b_lo + b_off[k] + kk * b_kk, hi_b, idesc);
                        }
                    }
                    umma_commit_e(barB_empty + 8 * stB);
                    umma_commit_e(barA_empty + 8 * stA);
                    if (++stB == stagesB) { stB = 0; phB ^= 1; }
                    if (++stA == SA) { stA = 0; phA ^= 1; }
                }
            } else
            for (int64_t t = t_beg; t < t_end; ++t) {
                mbar_wait(barA_full + 8 * stA, phA);
                tc_fence_after();
                const uint32_t aaddr = sA + stA * p.a_stage;
                for (int k = 0; k < ntap; ++k) {
                    mbar_wait(barB_full + 8 * stB, phB);
                    tc_fence_after();
                    const uint32_t baddr = sB + stB * p.b_stage;
                    const uint32_t tmem_d = tmem_base + k * p.TN;
                    const uint32_t a_lo = ((aaddr & 0x3FFFFu) >> 4) | ((((128 * rbA) >> 4) & 0x3FFFu) << 16);
                    const uint32_t b_lo = ((baddr & 0x3FFFFu) >> 4) | ((((128 * rbB) >> 4) & 0x3FFFu) << 16);
                    const uint32_t hi_a = ((8 * rbA) >> 4) | (1u << 14) | ((rbA == 128 ? 2u : 4u) << 29);
                    const uint32_t hi_b = ((8 * rbB) >> 4) | (1u << 14) | ((rbB == 128 ? 2u : 4u) << 29);
                    umma_bf16_lhr_e(tmem_d, a_lo, hi_a, b_lo, hi_b, idesc, t > t_beg ? 1u : 0u);
#pragma unroll
                    for (int kk = 1; kk < 8; ++kk)                         // 128 pixels = 8 x K16
                        umma_bf16_lh_e<1>(tmem_d, a_lo + kk * ((16 * rbA) >> 4), hi_a, b_lo + kk * ((16 * rbB) >> 4), hi_b, idesc);
                    umma_commit_e(barB_empty + 8 * stB);
                    if (++stB == p.stagesB) { stB = 0; phB ^= 1; }
                }
                umma_commit_e(barA_empty + 8 * stA);
                if (++stA == SA) { stA = 0; phA ^= 1; }
            }
            umma_commit_e(bar_done);
        }
    } else {
        const int quad = warp & 3, row = quad * 32 + lane;
        if (t_end > t_beg) {
            mbar_wait(bar_done, 0);
            tc_fence_after();
        }
        if (p.halo == 2) {
            const int g = row / p.chunkB, ci = row - g * p.chunkB;
            const int nq = p.kh * p.ncg;
            const int blk = p.rowN ? p.rowN : p.TNa;                 // columns of one (filter row, x-tap group) block
            for (int q = 0; q < nq; ++q) {
                // tap-stacked: q = r * ncg + cg; rows on N as well: q = cg * kh + j with block j holding filter row kh-1-j
                const int r = p.rowN ? p.kh - 1 - q % p.kh : q / p.ncg, cg = p.rowN ? q / p.kh : q % p.ncg;
                const int sx = cg * p.G + g;
                const bool ok = sx < p.kw && ci < p.Bp;
                float* out = partial + (((int64_t)split * taps + r * p.kw + sx) * p.Apad) * p.Bpad + ci;
                const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + q * blk;
                for (int c0 = 0; c0 < blk; c0 += 16) {
                    uint32_t v[16];
                    if (t_end > t_beg) { tmem_ld16(taddr + c0, v); tmem_ld_wait(); }
                    else {
#pragma unroll
                        for (int j = 0; j < 16; ++j) v[j] = 0u;
                    }
                    if (ok) {
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            if (c0 + j < p.Ap) out[(int64_t)(c0 + j) * p.Bpad] = __uint_as_float(v[j]);
                    }
                }
            }
        } else
        for (int k = 0; k < ntap; ++k) {
            float* out = partial + (((int64_t)split * taps + tap0 + k) * p.Apad + a0 + row) * p.Bpad + b0;
            const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + k * p.TN;
            for (int c0 = 0; c0 < p.TN; c0 += 16) {
                uint32_t v[16];
                if (t_end > t_beg) { tmem_ld16(taddr + c0, v); tmem_ld_wait(); }
                else {
#pragma unroll
                    for (int j = 0; j < 16; ++j) v[j] = 0u;                 // empty split: contributes zeros
                }
                if (a0 + row < p.Ap) {                                   // rows past the real channels are never read back
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        *reinterpret_cast<uint4*>(out + c0 + 4 * j) = make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)p.tmem_cols));
    }
}

int g_halo_mode = -1;   // -1: read MSX_TC_HALO from the environment (0 = streaming kernels only; measurement switch)

int halo_mode() {
    if (g_halo_mode < 0) {
        const char* e = getenv("MSX_TC_HALO");
        g_halo_mode = e ? atoi(e) : 1;
    }
    return g_halo_mode;
}

void pick_tile_pow2(int H, int W, int& tw, int& th) {
    double best = -1;
    tw = 16; th = 8;
    for (int w = 128; w >= 1; w >>= 1) {
        const int h = 128 / w;
        const double tiles = (double)((W + w - 1) / w) * ((H + h - 1) / h);
        const double eff = (double)W * H / (tiles * 128.0);
        if (eff > best + 1e-9) { best = eff; tw = w; th = h; }
    }
}

void wgrad_plan(const msx_conv_desc* d, WgParams& p) {
    // "small" side = the F-form destination (dy for Conv2d, x for ConvTranspose2d), "big" = the gathered side
    const int A = d->transposed ? d->cin : d->cout, B = d->transposed ? d->cout : d->cin;
    p.N = d->n;
    p.H = d->transposed ? d->h : d->oh; p.W = d->transposed ? d->w : d->ow;
    p.kh = d->kh; p.kw = d->kw; p.dil = d->dil; p.pad_h = d->pad_h; p.pad_w = d->pad_w;
    pick_tile_pow2(p.H, p.W, p.tw, p.th);
    p.tiles_x = (p.W + p.tw - 1) / p.tw;
    p.tiles_y = (p.H + p.th - 1) / p.th;
    p.pix_tiles = (int64_t)p.tiles_x * p.tiles_y * p.N;
    p.Ap = round_up(A, 8); p.Bp = round_up(B, 8);
    p.a_tiles = (p.Ap + 127) / 128;
    const int Bn = round_up(B, 16);
    p.b_tiles = (Bn + 255) / 256;
    p.TN = round_up((Bn + p.b_tiles - 1) / p.b_tiles, 16);
    p.chunkA = p.Ap <= 32 ? 32 : 64;
    p.chunkB = p.Bp <= 32 ? 32 : 64;
    p.groupsB = (p.TN + p.chunkB - 1) / p.chunkB;
    const int taps = p.kh * p.kw;
    p.tg = 512 / p.TN < taps ? 512 / p.TN : taps;
    p.tap_groups = (taps + p.tg - 1) / p.tg;
    int cols = 32;
    while (cols < p.tg * p.TN) cols <<= 1;
    p.tmem_cols = cols;
    p.Apad = p.a_tiles * 128; p.Bpad = p.b_tiles * p.TN;
    const int64_t units0 = (int64_t)p.a_tiles * p.b_tiles * p.tap_groups;
    // one CTA per SM and one wave: as many pixel splits as fit in 148 CTAs.  (Rounding UP gave 150-162 CTAs for the
    // 136- / 272- / 512-channel layers: a second wave of a dozen CTAs that doubled the kernel's time.)
    int64_t splits = units0 <= kNumSMs ? kNumSMs / units0 : 1;
    if (splits < 2) splits = (kNumSMs + units0 - 1) / units0;     // 75-148 units: two half-size CTAs per unit measured faster than one (544 ch: 45 vs 63 us)
    if (splits > p.pix_tiles) splits = p.pix_tiles;
    if (splits < 1) splits = 1;
    p.splits = (int)splits;
    p.a_stage = 128u * 256u;                                    // 128 channels x 128 pixels of bf16
    p.b_stage = (uint32_t)p.groupsB * 128u * (p.chunkB * 2);
    // halo mode (see conv_tc_halo_kernel): x channels fit one 64-wide group, 8x16 tiles cover the image well
    p.halo = 0; p.Wh = 0; p.Hh = 0;
    {
        const int span_w = (p.kw - 1) * p.dil, span_h = (p.kh - 1) * p.dil;
        const int tx = (p.W + 7) / 8, ty = (p.H + 15) / 16;
        const double eff = (double)p.W * p.H / ((double)tx * ty * 128.0);
        if (halo_mode() && d->stride == 1 && !d->transposed && p.Bp <= 64 && p.b_tiles == 1 && taps >= 2 && taps <= 9 && 8 + span_w <= 16 && eff >= 0.75) {
            p.halo = 1;
            p.tw = 8; p.th = 16; p.tiles_x = tx; p.tiles_y = ty;
            p.pix_tiles = (int64_t)tx * ty * p.N;
            if (splits > p.pix_tiles) { splits = p.pix_tiles; p.splits = (int)splits; }
            p.Wh = span_w > 0 ? 16 : 8;
            p.Hh = 16 + span_h;
            p.chunkB = 64; p.groupsB = 1;
            p.b_stage = ((uint32_t)p.Hh * p.Wh * 128u + 1023u) & ~1023u;
        }
    }
    p.TNa = 0; p.G = 0; p.ncg = 0; p.rowN = 0; p.HhA = 0; p.gs = p.dil; p.wide = 0;
    if (p.halo == 1 && p.Ap <= 128 && halo_mode() != 3) {
        // tap-stacked variant: see the halo == 2 branch of wgrad_tc_kernel
        const int chunkB = p.Bp <= 32 ? 32 : 64, G = 128 / chunkB, ncg = (p.kw + G - 1) / G;
        const int TNa = round_up(A, 16);
        const int gs = p.kw == 1 ? 0 : p.dil;
        const int Wh2 = (ncg * G - 1) * gs + 8;                       // exactly as wide as the stacked windows reach
        if (Wh2 <= 24 && p.kh * ncg * TNa <= 512 && p.kh * ncg <= 9) {
            p.halo = 2;
            p.chunkB = chunkB; p.G = G; p.ncg = ncg; p.TNa = TNa; p.gs = gs; p.Wh = Wh2;
            {
                static const int wide_on = [] { const char* e = getenv("MSX_WGRAD_WIDE"); return e ? atoi(e) : 1; }();
                const int txw = (p.W + 15) / 16, tyw = (p.H + 7) / 8;
                const double eff_t = (double)p.W * p.H / ((double)p.tiles_x * p.tiles_y * 128.0);
                const double eff_w = (double)p.W * p.H / ((double)txw * tyw * 128.0);
                if (wide_on && eff_w >= eff_t - 0.02) {
                    p.wide = 1;
                    p.tw = 16; p.th = 8; p.tiles_x = txw; p.tiles_y = tyw;
                    p.pix_tiles = (int64_t)txw * tyw * p.N;
                    p.Wh = Wh2 + 8;
                    p.Hh = 8 + (p.kh - 1) * p.dil;
                }
            }
            p.b_stage = ((uint32_t)p.Hh * p.Wh * (chunkB * 2) + 1023u) & ~1023u;
            p.a_tiles = 1; p.b_tiles = 1; p.tap_groups = 1; p.tg = taps;
            int64_t sp = kNumSMs < p.pix_tiles ? kNumSMs : p.pix_tiles;
            p.splits = (int)sp;
            int cols2 = 32;
            while (cols2 < p.kh * ncg * TNa) cols2 <<= 1;
            p.tmem_cols = cols2;
            // filter rows on the N axis too (see WgParams::rowN): the dy side must be ONE channel group
            static const int rows_on_n = [] { const char* e = getenv("MSX_WGRAD_ROWS_ON_N"); return e ? atoi(e) : 1; }();
            const int rowN = p.Ap <= 32 ? 32 : 64;
            if (rows_on_n && p.kh >= 2 && p.Ap <= rowN && p.kh * rowN <= 256 && ncg * p.kh * rowN <= 512) {
                p.rowN = rowN;
                p.chunkA = rowN;
                p.HhA = p.th + (p.kh - 1) * p.dil;
                p.Hh = p.th;                                            // x patch: x halo only
                p.b_stage = ((uint32_t)p.Hh * p.Wh * (chunkB * 2) + 1023u) & ~1023u;
                int cols3 = 32;
                while (cols3 < ncg * p.kh * rowN) cols3 <<= 1;
                p.tmem_cols = cols3;
            }
        }
    }
    // dy ring: 4 stages when a stage is small (few channels); the stage is as large as the boxes that fill it
    static const int sa_env = [] { const char* e = getenv("MSX_WGRAD_SA"); return e ? atoi(e) : 4; }();
    {
        const int groupsA = (((p.Ap < 128 ? p.Ap : 128) + p.chunkA - 1) / p.chunkA);
        const uint32_t need = (uint32_t)groupsA * 128u * (p.chunkA * 2);
        p.SA = (sa_env == 4 && need <= 16 * 1024) ? 4 : 2;
        if (p.SA == 4) p.a_stage = (need + 1023u) & ~1023u;
        p.a_box_bytes = 128u * (p.chunkA * 2);
        if (p.rowN) {                                                   // one dy box of HhA x 8 pixels
            p.a_box_bytes = (uint32_t)p.HhA * (uint32_t)p.tw * (p.chunkA * 2);
            p.a_stage = (p.a_box_bytes + 1023u) & ~1023u;
            p.SA = p.a_stage <= 24 * 1024 ? 4 : 2;
        }
    }
    int sb = (int)((kSmemBudget - p.SA * p.a_stage) / p.b_stage);
    p.stagesB = sb > kMaxStages ? kMaxStages : (sb < 2 ? 2 : sb);
#ifdef MSX_ABLATION
    { static const int dbg = [] { const char* e = getenv("MSX_DBG_WGRAD"); return e ? atoi(e) : 0; }(); p.dbg = dbg; }
#endif
}

}  // namespace

bool conv_tc_supported(const msx_conv_desc* d, bool fwd_pass) {
    (void)fwd_pass;
    if (d->dtype != MSX_BF16 || encode_fn() == nullptr) return false;
    if (d->kh * d->kw > kTapMaxTrain) return false;
    if (d->stride == 1) return !d->transposed && d->oh == d->h && d->ow == d->w;      // 'same' convs of the reference
    if (d->stride != 2 || d->dil != 1 || d->kh < 2 || d->kw < 2) return false;
    // stride 2 (Conv2d 3x3 p1 / 2x2 p0, ConvTranspose2d 3x3 p1 op1): fine side even and exactly twice the coarse side
    const int fh = d->transposed ? d->oh : d->h, fw = d->transposed ? d->ow : d->w;
    const int ch = d->transposed ? d->h : d->oh, cw = d->transposed ? d->w : d->ow;
    return fh % 2 == 0 && fw % 2 == 0 && fh == 2 * ch && fw == 2 * cw;
}

// Forward gather only: additionally the 7x7 stride-2 stem of a ResNet encoder (49 taps through the four phase views);
// its data / weight gradients are never needed (frozen teacher), so the training-side limit stays at 16 taps.
bool conv_tc_gather_supported(const msx_conv_desc* d, bool fwd_pass) {
    if (conv_tc_supported(d, fwd_pass)) return true;
    if (!fwd_pass || d->dtype != MSX_BF16 || encode_fn() == nullptr || d->transposed) return false;
    if (d->kh * d->kw > kTapMax || d->stride != 2 || d->dil != 1 || d->kh < 2 || d->kw < 2) return false;
    return d->h % 2 == 0 && d->w % 2 == 0 && d->h == 2 * d->oh && d->w == 2 * d->ow;
}

// 1x1 without padding: the weight gradient does not care where a pixel sits - walk the activations as ONE row of N*H*W pixels,
// so that a 128-pixel tile is one contiguous run for the TMA unit (see HaloParams::flat)
static const msx_conv_desc* flat_1x1(const msx_conv_desc* d, msx_conv_desc* tmp) {
    static const int on = [] { const char* e = getenv("MSX_FLAT_1X1"); return e ? atoi(e) : 1; }();
    const int64_t pixels = (int64_t)d->n * d->h * d->w;
    if (!on || d->kh != 1 || d->kw != 1 || d->stride != 1 || d->transposed || d->pad_h || d->pad_w || pixels >= (1ll << 31) - 128) return d;
    *tmp = *d;
    tmp->n = 1; tmp->h = tmp->oh = 1; tmp->w = tmp->ow = (int)pixels;
    return tmp;
}

size_t conv_wgrad_tc_ws_bytes(const msx_conv_desc* d) {
    msx_conv_desc flat_d;
    d = flat_1x1(d, &flat_d);
    WgParams p;
    wgrad_plan(d, p);
    return (size_t)p.splits * p.kh * p.kw * p.Apad * (size_t)p.Bpad * sizeof(float);
}

int wgrad_reduce_launch(const float* partial, int splits, int taps, int A, int B, int Apad, int Bpad, float* dw, cudaStream_t st);

// x: module input (big side, cin), dy: gradient of the module output (small side, cout); dw: [cout][cin][kh][kw] fp32.
static int make_src_maps(PhaseMaps& maps, const void* src, int Kp, int N, int Hs, int Ws, int lds, int stride,
                         int chunk, int tw, int th, unsigned used_mask);
static unsigned build_taps(const msx_conv_desc* d, bool f_form, int pa, int pb, TapTable& t);

int conv_wgrad_tc(const msx_conv_desc* d, const void* x, const void* dy, float* dw, void* ws, size_t ws_bytes, cudaStream_t st) {
    msx_conv_desc flat_d;
    d = flat_1x1(d, &flat_d);
    WgParams p;
    wgrad_plan(d, p);
    const size_t need = conv_wgrad_tc_ws_bytes(d);
    if (ws_bytes < need) return set_error(MSX_ERR_WORKSPACE, "conv wgrad (tcgen05): workspace %zu < %zu", ws_bytes, need);
    const void* small = d->transposed ? x : dy;
    const void* big = d->transposed ? dy : x;
    const int lda = d->transposed ? d->ldx : d->ldy, ldb = d->transposed ? d->ldy : d->ldx;
    const int Hb = d->transposed ? d->oh : d->h, Wb = d->transposed ? d->ow : d->w;
    const int A = d->transposed ? d->cin : d->cout, B = d->transposed ? d->cout : d->cin;
    const unsigned mask = build_taps(d, true, 0, 0, p.taps);       // big = small*stride + r*dil - pad
    if (mask == 0 || p.taps.n != p.kh * p.kw) return set_error(MSX_ERR_UNSUPPORTED, "conv wgrad (tcgen05): too many taps");
    CUtensorMap tmS;
    PhaseMaps tmG;
    {
        cuuint64_t dims[4] = {(cuuint64_t)p.Ap, (cuuint64_t)p.W, (cuuint64_t)p.H, (cuuint64_t)p.N};
        cuuint64_t strides[3] = {(cuuint64_t)lda * 2, (cuuint64_t)lda * 2 * p.W, (cuuint64_t)lda * 2 * p.W * p.H};
        cuuint32_t box[4] = {(cuuint32_t)p.chunkA, (cuuint32_t)p.tw, (cuuint32_t)(p.rowN ? p.HhA : p.th), 1};
        MSX_TRY(make_map(&tmS, 4, small, dims, strides, box, p.chunkA == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B));
    }
    MSX_TRY(make_src_maps(tmG, big, p.Bp, p.N, Hb, Wb, ldb, d->stride, p.chunkB, p.halo ? p.Wh : p.tw, p.halo ? p.Hh : p.th, mask));
    MSX_TRY(ensure_big_smem(reinterpret_cast<const void*>(wgrad_tc_kernel)));
    const size_t smem = 1024 + (size_t)p.SA * p.a_stage + (size_t)p.stagesB * p.b_stage + 64 + 16 * kMaxStages + 64;
    const int grid = p.a_tiles * p.b_tiles * p.tap_groups * p.splits;
    const int taps = p.kh * p.kw;
    const double Ms = (double)p.N * p.H * p.W, Mb = (double)p.N * Hb * Wb;
    {
        ProfScope ps__(K_WGRAD_TC, st, (Ms * A + Mb * B) * 2.0, 2.0 * Ms * A * B * taps);
        launch_pdl(wgrad_tc_kernel, dim3(grid), dim3(kThreads), smem, st, tmS, tmG, static_cast<float*>(ws), p);
    }
    MSX_LAUNCH_CHECK();
    return wgrad_reduce_launch(static_cast<const float*>(ws), p.splits, taps, A, B, p.Apad, p.Bpad, dw, st);
}

static inline uint32_t stat_smem_bytes(int ncp) { return 16u * (uint32_t)((ncp + 31) & ~31) * 8u + 16u; }

static int conv_gather_tc_halo(const msx_conv_desc* d, bool fwd_pass, const void* src, const void* packed, const ConvEpilogue& ep,
                               void* dst, cudaStream_t st, bool* used) {
    *used = false;
    const int mode = halo_mode();
    if (mode == 0) return MSX_OK;
    const int K = fwd_pass ? d->cin : d->cout, Nc = fwd_pass ? d->cout : d->cin;
    const int Kp = round_up(K, 8), Np = round_up(Nc, 16), taps = d->kh * d->kw;
    // 64 < Kp <= 80 (DUCK-Net's 68 channels): the last 16 channels travel as a second, 32 B-per-pixel patch (SWIZZLE_32B)
    static const int halo_tail_on = [] { const char* e = getenv("MSX_HALO_TAIL"); return e ? atoi(e) : 1; }();
    const bool tail = Kp > 64 && Kp <= 80 && halo_tail_on;
    if ((Kp > 64 && !tail) || Np > 256 || taps > 9) return MSX_OK;         // taps == 1 (1x1): no halo, but resident weights + lean issue loop
    if ((size_t)taps * Np * (tail ? 160 : 128) > (tail ? 120u : 80u) * 1024) return MSX_OK;
    const int span_w = (d->kw - 1) * d->dil, span_h = (d->kh - 1) * d->dil;
    if (8 + span_w > 16) return MSX_OK;
    HaloParams p;
    p.N = d->n; p.H = d->h; p.W = d->w;
    p.tiles_x = (p.W + 7) / 8; p.tiles_y = (p.H + 15) / 16;
    static const int flat_on = [] { const char* e = getenv("MSX_FLAT_1X1"); return e ? atoi(e) : 1; }();
    const int64_t pixels = (int64_t)d->n * d->h * d->w;
    p.flat = (flat_on && taps == 1 && d->pad_h == 0 && d->pad_w == 0 && pixels < (1ll << 31) - 128) ? 1 : 0;
    if (p.flat) { p.N = 1; p.H = 1; p.W = (int)pixels; p.tiles_x = (int)((pixels + 127) / 128); p.tiles_y = 1; }
    const double eff = p.flat ? 1.0 : (double)p.W * p.H / ((double)p.tiles_x * p.tiles_y * 128.0);
    if (eff < 0.75) return MSX_OK;
    p.total_tiles = p.tiles_x * p.tiles_y * p.N;
    p.kh = d->kh; p.kw = d->kw;
    if (fwd_pass) { p.mul_t = d->dil; p.off_h = -d->pad_h; p.off_w = -d->pad_w; p.oy_min = p.off_h; p.ox_min = p.off_w; }
    else          { p.mul_t = -d->dil; p.off_h = d->pad_h; p.off_w = d->pad_w; p.oy_min = p.off_h - span_h; p.ox_min = p.off_w - span_w; }
    // Halo exactly as wide as the taps reach (MSX_HALO_TIGHT=0: padded to 16 pixels).  The 128B swizzle follows the
    // absolute shared-memory address, so the row pitch need not be a multiple of the 1024 B swizzle period.
    static const int tight = [] { const char* e = getenv("MSX_HALO_TIGHT"); return e ? atoi(e) : 1; }();
    p.Hh = 16 + span_h;
    p.BN = Np; p.ksteps = tail ? 4 : (Kp + 15) / 16;
    p.Nc = Nc; p.ncp = round_up(Nc, 8); p.ldd = fwd_pass ? d->ldy : d->ldx;
    static const int rows64 = [] { const char* e = getenv("MSX_HALO_ROWS64"); return e ? atoi(e) : 1; }();
    p.rb = (Kp <= 32 && rows64) ? 64u : 128u;
    p.b_tap_bytes = (uint32_t)p.BN * p.rb;
    p.b_tail_bytes = tail ? (uint32_t)p.BN * 32 : 0;
    p.b_tail_base = ((uint32_t)taps * p.b_tap_bytes + 1023u) & ~1023u;
    const uint32_t stat_bytes = ep.stats ? stat_smem_bytes(p.ncp) : 0;
    const uint32_t budget = tail ? 214u * 1024u : kSmemBudget;     // 115 KB of resident weights: three stages only just fit
    const uint32_t bres = ((p.b_tail_base + (uint32_t)taps * p.b_tail_bytes + 1023u) & ~1023u) + stat_bytes;      // resident bytes: weights + statistics
    p.nacc = 512 / p.BN < kMaxAcc ? 512 / p.BN : kMaxAcc;
    if (p.nacc % 2) --p.nacc;                                     // two epilogue groups alternate tiles
#ifdef MSX_ABLATION
    { static const int dbg = [] { const char* e = getenv("MSX_DBG_HALO"); return e ? atoi(e) : 0; }(); p.dbg = dbg; }
#endif
    static const int iss = [] { const char* e = getenv("MSX_HALO_ISSUERS"); return e && atoi(e) == 1 ? 1 : 2; }();
    static const int want_pair = [] { const char* e = getenv("MSX_HALO_PAIR"); return e ? atoi(e) : 0; }();
    p.issuers = p.nacc >= 2 ? iss : 1;
    // Pair mode (MSX_HALO_PAIR=1, off by default): one TMA box for two tiles side by side (16 x 16 pixels + halo), issuer
    // and epilogue group j taking tile j.  Measured: 10 % faster for wide halos (3x3 d2 101 -> 89 us, 1x7 92 -> 81 us at
    // 17 channels, 352x352), nothing for 3x3 d1, 10 % slower at 64 channels (fewer stages fit); the whole step is unchanged,
    // so the per-box cost of the TMA unit is not what bounds these layers.  Kept for the next round's layout work.
    p.boxes_x = (p.tiles_x + 1) / 2;
    p.total_boxes = p.boxes_x * p.tiles_y * p.N;
    p.tw = (want_pair && !p.flat && !tail && tight && p.issuers == 2 && (p.tiles_x % 2 == 0 || p.tiles_x >= 9)) ? 2 : 1;
    for (;;) {
        p.Wh = p.tw == 2 ? 16 + span_w : (tight ? 8 + span_w : (span_w > 0 ? 16 : 8));
        p.a_bytes = (uint32_t)p.Hh * p.Wh * p.rb;
        p.a_tail_off = (p.a_bytes + 1023u) & ~1023u;
        p.a_bytes_tail = tail ? (uint32_t)p.Hh * p.Wh * 32 : 0;
        p.a_stage = (p.a_tail_off + p.a_bytes_tail + 1023u) & ~1023u;
        int stages = bres + 2 * p.a_stage <= budget ? (int)((budget - bres) / p.a_stage) : 0;
        p.stages = stages > kMaxStages ? kMaxStages : stages;
        if (p.stages >= 2 || p.tw == 1) break;
        p.tw = 1;                                                 // the wide box does not leave room for two stages
    }
    if (p.stages < 2) return MSX_OK;
#ifdef MSX_ABLATION
    {
        static const int st = [] { const char* e = getenv("MSX_DBG_STAGES"); return e ? atoi(e) : 0; }();
        static const int na = [] { const char* e = getenv("MSX_DBG_NACC"); return e ? atoi(e) : 0; }();
        if (st >= 2 && st < p.stages) p.stages = st;
        if (na >= 2 && na < p.nacc) p.nacc = na & ~1;
    }
#endif
    // Two issuers need an even stage (and accumulator) count: then every stage / accumulator belongs to ONE issuer and
    // its mbarrier phases advance only with that thread's own progress.  With an odd count a stage alternates between
    // the issuers, one of them can run two phases ahead of the other and a parity wait passes on the wrong phase
    // (seen as a launch failure in the full step with 3 stages; the unit tests did not catch it).
    if (p.issuers == 2 && p.tw == 1) p.stages &= ~1;               // (pair mode: both issuers share every stage, released by both)
    int cols = 32;
    while (cols < p.nacc * p.BN) cols <<= 1;
    p.tmem_cols = cols;
    const int lds = fwd_pass ? d->ldx : d->ldy;
    CUtensorMap tmA, tmB, tmAt, tmBt;
    {
        cuuint64_t dims[4] = {(cuuint64_t)Kp, (cuuint64_t)p.W, (cuuint64_t)p.H, (cuuint64_t)p.N};
        cuuint64_t strides[3] = {(cuuint64_t)lds * 2, (cuuint64_t)lds * 2 * p.W, (cuuint64_t)lds * 2 * p.W * p.H};
        const cuuint32_t bw = p.flat ? 128u : (cuuint32_t)p.Wh, bh = p.flat ? 1u : (cuuint32_t)p.Hh;      // flat: Wh x Hh = 8 x 16 rows as one run
        cuuint32_t box[4] = {p.rb / 2, bw, bh, 1};
        MSX_TRY(make_map(&tmA, 4, src, dims, strides, box, p.rb == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B));
        tmAt = tmA;
        if (tail) {
            cuuint32_t boxt[4] = {16, bw, bh, 1};
            MSX_TRY(make_map(&tmAt, 4, src, dims, strides, boxt, CU_TENSOR_MAP_SWIZZLE_32B));
        }
    }
    {
        cuuint64_t dims[3] = {(cuuint64_t)Kp, (cuuint64_t)Np, (cuuint64_t)taps};
        cuuint64_t strides[2] = {(cuuint64_t)Kp * 2, (cuuint64_t)Kp * 2 * Np};
        cuuint32_t box[3] = {p.rb / 2, (cuuint32_t)p.BN, 1};
        MSX_TRY(make_map(&tmB, 3, packed, dims, strides, box, p.rb == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B));
        tmBt = tmB;
        if (tail) {
            cuuint32_t boxt[3] = {16, (cuuint32_t)p.BN, 1};
            MSX_TRY(make_map(&tmBt, 3, packed, dims, strides, boxt, CU_TENSOR_MAP_SWIZZLE_32B));
        }
    }
    const size_t smem = 1024 + bres + (size_t)p.stages * p.a_stage + 16 * kMaxStages + 16 * kMaxAcc + 96;
    const int units = p.tw == 2 ? p.total_boxes : p.total_tiles;
    const int grid = units < kNumSMs ? units : kNumSMs;
    const double M = (double)p.N * p.H * p.W;
    typedef void (*HaloKernel)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, bf16*, const ConvEpilogue, const HaloParams);
    HaloKernel kern = nullptr;
#define MSX_HALO_CASE(T, K) if (taps == T && p.ksteps == K && !tail) kern = conv_tc_halo_kernel<T, K, false>;
    MSX_HALO_CASE(9, 1) MSX_HALO_CASE(9, 2) MSX_HALO_CASE(9, 3) MSX_HALO_CASE(9, 4)
    MSX_HALO_CASE(7, 1) MSX_HALO_CASE(7, 2) MSX_HALO_CASE(7, 3) MSX_HALO_CASE(7, 4)
    MSX_HALO_CASE(1, 1) MSX_HALO_CASE(1, 2) MSX_HALO_CASE(1, 3) MSX_HALO_CASE(1, 4)
#undef MSX_HALO_CASE
    if (tail && taps == 9) kern = conv_tc_halo_kernel<9, 4, true>;
    if (tail && taps == 7) kern = conv_tc_halo_kernel<7, 4, true>;
    if (tail && taps == 1) kern = conv_tc_halo_kernel<1, 4, true>;
    if (!kern) return MSX_OK;                                     // other tap counts: streaming kernel
    MSX_TRY(ensure_big_smem(reinterpret_cast<const void*>(kern)));     // > 48 KB dynamic smem: once per device and instantiation
    {
        ProfScope ps__(K_CONV_TC, st, M * (K + Nc) * 2.0, 2.0 * M * Nc * K * taps);
        launch_pdl(kern, dim3(grid), dim3(kHaloThreads2), smem, st, tmA, tmB, tmAt, tmBt, static_cast<bf16*>(dst), ep, p);
    }
    MSX_LAUNCH_CHECK();
    if (ep.stats && !ep.bnf.ticket) MSX_TRY(bn_finalize_launch(ep.bnf, grid, st));
    note_conv_engine(MSX_ENGINE_TCGEN05);
    *used = true;
    return MSX_OK;
}

// Source activation seen through tensor maps: one map for stride 1, or the four stride-2 phase views
// (phase (py, px) holds pixels (2y + py, 2x + px)); a strided convolution then reads every tap as a plain
// shifted box of one phase, so zero padding stays TMA out-of-bounds fill.
static int make_src_maps(PhaseMaps& maps, const void* src, int Kp, int N, int Hs, int Ws, int lds, int stride,
                         int chunk, int tw, int th, unsigned used_mask) {
    const CUtensorMapSwizzle sw = chunk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : (chunk == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    cuuint32_t box[4] = {(cuuint32_t)chunk, (cuuint32_t)tw, (cuuint32_t)th, 1};
    if (stride == 1) {
        cuuint64_t dims[4] = {(cuuint64_t)Kp, (cuuint64_t)Ws, (cuuint64_t)Hs, (cuuint64_t)N};
        cuuint64_t strides[3] = {(cuuint64_t)lds * 2, (cuuint64_t)lds * 2 * Ws, (cuuint64_t)lds * 2 * Ws * Hs};
        MSX_TRY(make_map(&maps.m[0], 4, src, dims, strides, box, sw));
        maps.m[1] = maps.m[2] = maps.m[3] = maps.m[0];
        return MSX_OK;
    }
    int first = -1;
    for (int ph = 0; ph < 4; ++ph) {
        if (!(used_mask & (1u << ph))) continue;
        const int py = ph >> 1, px = ph & 1;
        const char* base = static_cast<const char*>(src) + ((size_t)py * Ws + px) * lds * 2;
        cuuint64_t dims[4] = {(cuuint64_t)Kp, (cuuint64_t)((Ws - px + 1) / 2), (cuuint64_t)((Hs - py + 1) / 2), (cuuint64_t)N};
        cuuint64_t strides[3] = {(cuuint64_t)lds * 4, (cuuint64_t)lds * 4 * Ws, (cuuint64_t)lds * 2 * Ws * Hs};
        MSX_TRY(make_map(&maps.m[ph], 4, base, dims, strides, box, sw));
        if (first < 0) first = ph;
    }
    for (int ph = 0; ph < 4; ++ph)
        if (!(used_mask & (1u << ph))) maps.m[ph] = maps.m[first];
    return MSX_OK;
}

static inline int mod2(int v) { return ((v % 2) + 2) % 2; }

// Tap table of one gather launch.  f_form: src = dst*stride + r*dil - pad.  Otherwise (b-form)
// src = (dst + pad - r*dil) / stride, and for stride 2 the launch covers destination phase (pa, pb).
static unsigned build_taps(const msx_conv_desc* d, bool f_form, int pa, int pb, TapTable& t) {
    unsigned mask = 0;
    t.n = 0;
    for (int r = 0; r < d->kh; ++r)
        for (int s = 0; s < d->kw; ++s) {
            int map = 0, dy, dx;
            if (f_form) {
                const int vy = r * d->dil - d->pad_h, vx = s * d->dil - d->pad_w;
                if (d->stride == 1) { dy = vy; dx = vx; }
                else { const int py = mod2(vy), px = mod2(vx); map = py * 2 + px; dy = (vy - py) / 2; dx = (vx - px) / 2; }
            } else {
                const int vy = pa + d->pad_h - r * d->dil, vx = pb + d->pad_w - s * d->dil;
                if (d->stride == 1) { dy = vy; dx = vx; }
                else { if (mod2(vy) || mod2(vx)) continue; dy = (vy - mod2(vy)) / 2; dx = (vx - mod2(vx)) / 2; }
            }
            if (t.n >= kTapMax) return 0;
            t.map[t.n] = (signed char)map; t.dy[t.n] = (signed char)dy; t.dx[t.n] = (signed char)dx;
            t.wtap[t.n] = (signed char)(r * d->kw + s);
            mask |= 1u << map;
            ++t.n;
        }
    return mask;
}

// stats_fused (nullable): set when ep.stats was honoured (the kernel produced and finalized the BatchNorm statistics);
// geometries that cannot (several launches per convolution, more than one tile of output channels) leave it false and
// the caller runs the stand-alone statistics pass.
int conv_gather_tc(const msx_conv_desc* d, bool fwd_pass, const void* src, const void* packed, const ConvEpilogue& ep_in,
                   void* dst, cudaStream_t st, bool* stats_fused) {
    const bool f_form = d->transposed ? !fwd_pass : fwd_pass;
    if (stats_fused) *stats_fused = false;
    ConvEpilogue ep = ep_in;
    if (d->stride == 1 && !d->transposed) {
        bool used = false;
        MSX_TRY(conv_gather_tc_halo(d, fwd_pass, src, packed, ep, dst, st, &used));
        if (used) {
            if (stats_fused) *stats_fused = ep.stats != 0;
            return MSX_OK;
        }
    }
    TcParams p;
    const int K = fwd_pass ? d->cin : d->cout, Nc = fwd_pass ? d->cout : d->cin;
    const int lds = fwd_pass ? d->ldx : d->ldy;
    const int Hs = fwd_pass ? d->h : d->oh, Ws = fwd_pass ? d->w : d->ow;
    p.Hd = fwd_pass ? d->oh : d->h; p.Wd = fwd_pass ? d->ow : d->w;
    p.N = d->n;
    p.Kp = round_up(K, 8);
    p.Nc = Nc; p.ncp = round_up(Nc, 8); p.ldd = fwd_pass ? d->ldy : d->ldx;
    const int Np = round_up(Nc, 16);
    p.n_tiles = (Np + 255) / 256;
    {
        // Output-channel tiling by a small cost model: rounds of the persistent grid x time of one K-step, which is
        // bound by the operand fetch (~32 cycles for the 128-row A tile + N/4 for B).  Few-pixel layers (11x11, 22x22) fill
        // more SMs with narrower N tiles: 544 -> 544 at 11x11 is 16 x 3 = 48 tiles of N = 192 by default, 144 tiles of N = 64
        // here; 272 channels at 22x22 go from 160 tiles (two rounds for 12 stragglers) to 240 narrower ones.
        const bool four_phase = (d->transposed ? fwd_pass : !fwd_pass) && d->stride == 2;
        const int gh = four_phase ? (p.Hd + 1) / 2 : p.Hd, gw = four_phase ? (p.Wd + 1) / 2 : p.Wd;
        int tw0, th0;
        pick_tile(gh, gw, tw0, th0);
        const int64_t m_tiles = (int64_t)((gw + tw0 - 1) / tw0) * ((gh + th0 - 1) / th0) * p.N;
        double best = 1e30;
        int best_n = p.n_tiles;
        for (int nt = p.n_tiles; nt <= p.n_tiles + 8 && nt * 16 <= Np; ++nt) {
            const int bn = round_up((Np + nt - 1) / nt, 16);
            if ((bn * (nt - 1)) >= Np) continue;                 // last tile would be empty
            const double rounds = (double)((m_tiles * nt + kNumSMs - 1) / kNumSMs);
            const double cost = rounds * (32.0 + bn / 4.0);
            if (cost < best * 0.97) { best = cost; best_n = nt; }   // prefer fewer tiles unless clearly better
        }
        p.n_tiles = best_n;
    }
    p.BN = round_up((Np + p.n_tiles - 1) / p.n_tiles, 16);
    p.chunk = p.Kp <= 32 ? 32 : 64;
    p.kchunks = (p.Kp + p.chunk - 1) / p.chunk;
    const uint32_t row_bytes = p.chunk * 2;
    p.b_bytes = (uint32_t)p.BN * row_bytes;
    {
        static const int tail_on = [] { const char* e = getenv("MSX_TC_TAIL"); return e ? atoi(e) : 1; }();
        const int rem = p.Kp % 64;                               // DUCK-Net's 17 * 2^k widths: 72 -> 8, 136 -> 8, 272 -> 16, 544 -> 32
        p.tail = (tail_on && p.chunk == 64 && p.kchunks >= 2 && rem > 0 && rem <= 32) ? (rem <= 16 ? 16 : 32) : 0;
        p.b_bytes_tail = (uint32_t)p.BN * p.tail * 2;
    }
    p.a_stage = 128 * row_bytes;                                 // full 128-row operand tile
    p.b_stage = (p.b_bytes + 1023u) & ~1023u;
    const int nphase = (!f_form && d->stride == 2) ? 4 : 1;      // b-form stride 2: one launch per destination phase
    // several tiles of output channels: a CTA's statistics rows cover ALL columns (it may visit several n tiles; the columns it
    // never sees stay zero), which costs shared memory - up to 72 KB (544 channels) are traded for the stand-alone pass
    static const int stats_ntiles = [] { const char* e = getenv("MSX_STATS_NTILES"); return e ? atoi(e) : 1; }();
    if (nphase != 1 || (p.n_tiles != 1 && (!stats_ntiles || stat_smem_bytes(p.ncp) > 72u * 1024u))) ep.stats = 0;
    const uint32_t stat_bytes = ep.stats ? stat_smem_bytes(p.ncp) : 0;
    int stages = (int)((kSmemBudget - stat_bytes) / (p.a_stage + p.b_stage));
    p.stages = stages > kMaxStages ? kMaxStages : (stages < 2 ? 2 : stages);
    int cols = 32;
    while (cols < 2 * p.BN) cols <<= 1;
    p.tmem_cols = cols;
    const size_t smem = 1024 + (size_t)p.stages * (p.a_stage + p.b_stage) + 16 * kMaxStages + 64 + stat_bytes;
    const CUtensorMapSwizzle sw = p.chunk == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
    CUtensorMap tmB, tmB_tail;
    {
        cuuint64_t dims[3] = {(cuuint64_t)p.Kp, (cuuint64_t)Np, (cuuint64_t)(d->kh * d->kw)};
        cuuint64_t strides[2] = {(cuuint64_t)p.Kp * 2, (cuuint64_t)p.Kp * 2 * Np};
        cuuint32_t box[3] = {(cuuint32_t)p.chunk, (cuuint32_t)p.BN, 1};
        MSX_TRY(make_map(&tmB, 3, packed, dims, strides, box, sw));
        tmB_tail = tmB;
        if (p.tail) {
            cuuint32_t boxt[3] = {(cuuint32_t)p.tail, (cuuint32_t)p.BN, 1};
            MSX_TRY(make_map(&tmB_tail, 3, packed, dims, strides, boxt, p.tail == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B));
        }
    }
    MSX_TRY(ensure_big_smem(reinterpret_cast<const void*>(conv_tc_kernel)));
    for (int ph = 0; ph < nphase; ++ph) {
        const int pa = ph >> 1, pb = ph & 1;
        const unsigned mask = build_taps(d, f_form, pa, pb, p.taps);
        if (p.taps.n == 0 || mask == 0) return set_error(MSX_ERR_UNSUPPORTED, "conv (tcgen05): empty tap set for this geometry");
        if (nphase == 4) { p.dmul = 2; p.da = pa; p.db = pb; p.H = (p.Hd - pa + 1) / 2; p.W = (p.Wd - pb + 1) / 2; }
        else             { p.dmul = 1; p.da = 0; p.db = 0; p.H = p.Hd; p.W = p.Wd; }
        pick_tile(p.H, p.W, p.tw, p.th);
        p.tiles_x = (p.W + p.tw - 1) / p.tw;
        p.tiles_y = (p.H + p.th - 1) / p.th;
        p.total_tiles = p.tiles_x * p.tiles_y * p.N * p.n_tiles;
        p.a_bytes = (uint32_t)p.tw * p.th * row_bytes;
        p.a_bytes_tail = (uint32_t)p.tw * p.th * p.tail * 2;
        PhaseMaps maps, maps_tail;
        MSX_TRY(make_src_maps(maps, src, p.Kp, p.N, Hs, Ws, lds, f_form ? d->stride : 1, p.chunk, p.tw, p.th, mask));
        if (p.tail) MSX_TRY(make_src_maps(maps_tail, src, p.Kp, p.N, Hs, Ws, lds, f_form ? d->stride : 1, p.tail, p.tw, p.th, mask));
        else maps_tail = maps;
        const int grid = p.total_tiles < kNumSMs ? p.total_tiles : kNumSMs;
        const double Mg = (double)p.N * p.H * p.W, Ms = (double)p.N * Hs * Ws / nphase;
        {
            ProfScope ps__(K_CONV_TC, st, (Mg * Nc + Ms * K) * 2.0, 2.0 * Mg * Nc * K * p.taps.n);
            launch_pdl(conv_tc_kernel, dim3(grid), dim3(kHaloThreads), smem, st, maps, tmB, maps_tail, tmB_tail, static_cast<bf16*>(dst), ep, p);
        }
        MSX_LAUNCH_CHECK();
        if (ep.stats && !ep.bnf.ticket) MSX_TRY(bn_finalize_launch(ep.bnf, grid, st));
    }
    note_conv_engine(MSX_ENGINE_TCGEN05);
    if (stats_fused) *stats_fused = ep.stats != 0;
    return MSX_OK;
}

}  // namespace msx
