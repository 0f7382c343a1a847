// How fast does one SM's TMA unit fill shared memory?  Tensor-map boxes with 32- / 64- / 128-byte rows against a flat
// cp.async.bulk copy of the same bytes, 148 persistent CTAs, a 4-stage ring whose consumer only frees the stage.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/micro/tma_rate tools/micro/tma_rate.cu -lcuda
// The numbers are quoted in DESIGN.md section 5 (they decide the layout of the packed conv weights).
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t a, uint32_t n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(a), "r"(n)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t a, uint32_t b) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(a), "r"(b) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t a) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(a) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t a, uint32_t ph) {
    asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(a), "r"(ph) : "memory");
}
__device__ __forceinline__ void tma_2d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(m), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_4d(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
                 ::"r"(dst), "l"(m), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void bulk_1d(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

constexpr int kStages = 4;
__global__ void __launch_bounds__(64, 1) rate_kernel(const __grid_constant__ CUtensorMap tm, const uint8_t* src, int mode, int rows_per_box,
                                                     uint32_t stage_bytes, int iters, int64_t span_rows, int row_bytes, long long* cycles) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const uint32_t bars = base + kStages * stage_bytes;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(bars + 8 * s, 1); mbar_init(bars + 64 + 8 * s, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const long long t0 = clock64();
    if (threadIdx.x == 0) {
        int st = 0; uint32_t ph = 0;
        int64_t row = (int64_t)blockIdx.x * rows_per_box;
        for (int i = 0; i < iters; ++i) {
            mbar_wait(bars + 64 + 8 * st, ph ^ 1);
            mbar_expect_tx(bars + 8 * st, stage_bytes);
            if (mode == 0) tma_2d(base + st * stage_bytes, &tm, bars + 8 * st, 0, (int)row);
            else bulk_1d(base + st * stage_bytes, src + row * row_bytes, stage_bytes, bars + 8 * st);
            row += (int64_t)gridDim.x * rows_per_box;
            if (row + rows_per_box > span_rows) row = (int64_t)blockIdx.x * rows_per_box;
            if (++st == kStages) { st = 0; ph ^= 1; }
        }
    } else if (threadIdx.x == 32) {
        int st = 0; uint32_t ph = 0;
        for (int i = 0; i < iters; ++i) {
            mbar_wait(bars + 8 * st, ph);
            mbar_arrive(bars + 64 + 8 * st);
            if (++st == kStages) { st = 0; ph ^= 1; }
        }
        cycles[blockIdx.x] = clock64() - t0;
    }
}

// The halo conv kernel's access pattern: NHWC activation [16][352][352][ld] bf16, one box of {box_c, bw, bh} per 8 x 16
// pixel tile (bw = 8 + x halo, bh = 16 + y halo), tiles walked like the persistent kernel walks them.
__global__ void __launch_bounds__(64, 1) conv_pattern_kernel(const __grid_constant__ CUtensorMap tm, uint32_t stage_bytes, int tiles_x, int tiles_y,
                                                             int total, long long* cycles, int kStages) {
    extern __shared__ uint8_t raw[];
    const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
    const uint32_t stage_pitch = (stage_bytes + 1023u) & ~1023u;
    const uint32_t bars = base + kStages * stage_pitch;
    if (threadIdx.x == 0) {
        for (int s = 0; s < kStages; ++s) { mbar_init(bars + 8 * s, 1); mbar_init(bars + 128 + 8 * s, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const long long t0 = clock64();
    if (threadIdx.x == 0) {
        int st = 0; uint32_t ph = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x) {
            const int img = t / (tiles_x * tiles_y), rem = t % (tiles_x * tiles_y);
            mbar_wait(bars + 128 + 8 * st, ph ^ 1);
            mbar_expect_tx(bars + 8 * st, stage_bytes);
            tma_4d(base + st * stage_pitch, &tm, bars + 8 * st, 0, (rem % tiles_x) * 8 - 1, (rem / tiles_x) * 16 - 1, img);
            if (++st == kStages) { st = 0; ph ^= 1; }
        }
    } else if (threadIdx.x == 32) {
        int st = 0; uint32_t ph = 0;
        for (int t = blockIdx.x; t < total; t += gridDim.x) {
            mbar_wait(bars + 8 * st, ph);
            mbar_arrive(bars + 128 + 8 * st);
            if (++st == kStages) { st = 0; ph ^= 1; }
        }
        cycles[blockIdx.x] = clock64() - t0;
    }
}

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
    void* f = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q);
    EncodeFn enc = (EncodeFn)f;
    const size_t total = 1ull << 30;
    uint8_t* buf; cudaMalloc(&buf, total); cudaMemset(buf, 1, total);
    long long* cyc; cudaMalloc(&cyc, 148 * 8);
    cudaFuncSetAttribute(rate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    printf("%-34s %10s %10s %12s\n", "load", "span", "us", "B/clk/SM");
    const int grids[2] = {148, 48};
    for (int gi = 0; gi < 2; ++gi)
    for (int span_mb : {32, 1024})                       // 32 MB: L2 resident after the first pass; 1 GB: from HBM
    for (int cfg = 0; cfg < 6; ++cfg) {
        // cfg: row bytes / rows per box / mode
        const int row_bytes = cfg == 0 ? 128 : cfg == 1 ? 64 : cfg == 2 ? 32 : cfg == 3 ? 128 : cfg == 4 ? 128 : 128;
        const int rows = cfg == 3 ? 256 : cfg == 4 ? 64 : 128;
        const int mode = cfg == 5 ? 1 : 0;
        const uint32_t stage_bytes = (uint32_t)rows * row_bytes;
        const int64_t span_rows = (int64_t)span_mb * 1024 * 1024 / row_bytes;
        CUtensorMap tm;
        cuuint64_t dims[2] = {(cuuint64_t)row_bytes / 2, (cuuint64_t)span_rows};
        cuuint64_t strides[1] = {(cuuint64_t)row_bytes};
        cuuint32_t box[2] = {(cuuint32_t)row_bytes / 2, (cuuint32_t)rows}, es[2] = {1, 1};
        CUtensorMapSwizzle sw = row_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : row_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
        if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, buf, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); return 1; }
        const int iters = 4000, grid = grids[gi];
        const size_t smem = 1024 + kStages * stage_bytes + 256;
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            rate_kernel<<<grid, 64, smem>>>(tm, buf, mode, rows, stage_bytes, iters, span_rows, row_bytes, cyc);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        if (cudaGetLastError() != cudaSuccess) { printf("launch failed\n"); return 1; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long h[148]; cudaMemcpy(h, cyc, grid * 8, cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < grid; ++i) avg += (double)h[i]; avg /= grid;
        char name[64];
        snprintf(name, sizeof name, "%s %d x %d B, %d CTAs", mode ? "bulk " : "tensor", rows, row_bytes, grid);
        printf("%-34s %7d MB %10.1f %12.1f\n", name, span_mb, ms * 1e3, (double)iters * stage_bytes / avg);
    }
    // conv access pattern, 16 x 352 x 352 pixels
    printf("\n%-44s %10s %12s %14s\n", "NHWC box per 8x16 tile (148 CTAs)", "us", "cyc/tile", "read GB/s");
    cudaFuncSetAttribute(conv_pattern_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    struct Pat { int ld, box_c, bw, bh; const char* what; } pats[] = {
        {24, 32, 8, 16, "ld 24, box 32ch x 8 x 16 (1x1, 64 B rows)"},   {24, 64, 8, 16, "ld 24, box 64ch x 8 x 16 (1x1, 128 B rows)"},
        {24, 32, 10, 18, "ld 24, box 32ch x 10 x 18 (3x3, 64 B rows)"}, {24, 64, 10, 18, "ld 24, box 64ch x 10 x 18 (3x3, 128 B rows)"},
        {32, 32, 10, 18, "ld 32, box 32ch x 10 x 18 (3x3, 64 B rows)"}, {64, 64, 10, 18, "ld 64, box 64ch x 10 x 18 (3x3, 128 B rows)"},
        {24, 32, 14, 16, "ld 24, box 32ch x 14 x 16 (1x7)"},            {24, 32, 12, 20, "ld 24, box 32ch x 12 x 20 (3x3 d2)"},
    };
    for (int stages : {4, 8, 16})
    for (const Pat& pt : pats) {
        CUtensorMap tm;
        cuuint64_t dims[4] = {(cuuint64_t)pt.ld, 352, 352, 16};
        cuuint64_t strides[3] = {(cuuint64_t)pt.ld * 2, (cuuint64_t)pt.ld * 2 * 352, (cuuint64_t)pt.ld * 2 * 352 * 352};
        cuuint32_t box[4] = {(cuuint32_t)pt.box_c, (cuuint32_t)pt.bw, (cuuint32_t)pt.bh, 1}, es[4] = {1, 1, 1, 1};
        CUtensorMapSwizzle sw = pt.box_c == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
        if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, buf, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); return 1; }
        const uint32_t stage_bytes = (uint32_t)pt.box_c * 2 * pt.bw * pt.bh;
        const int tiles_x = 44, tiles_y = 22, total_tiles = tiles_x * tiles_y * 16;
        const size_t smem = 1024 + stages * ((stage_bytes + 1023) & ~1023u) + 512;
        if (smem > 200 * 1024) continue;
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 2; ++rep) {
            cudaMemset(buf + (512u << 20), 2, 256u << 20);            // flush L2
            cudaEventRecord(e0);
            conv_pattern_kernel<<<148, 64, smem>>>(tm, stage_bytes, tiles_x, tiles_y, total_tiles, cyc, stages);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
        }
        if (cudaGetLastError() != cudaSuccess) { printf("launch failed\n"); return 1; }
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        long long h[148]; cudaMemcpy(h, cyc, 148 * 8, cudaMemcpyDeviceToHost);
        double avg = 0; for (int i = 0; i < 148; ++i) avg += (double)h[i]; avg /= 148;
        printf("%2d stages  %-44s %10.1f %12.1f %14.0f\n", stages, pt.what, ms * 1e3, avg / (total_tiles / 148.0), 16.0 * 352 * 352 * pt.ld * 2 / (ms * 1e6));
    }
    return 0;
}
