// Error channel and misc. entry points of libmsx.so.
#include <stdarg.h>

#include <atomic>
#include <mutex>
#include <vector>

#include "common.cuh"

namespace msx {
static thread_local char g_err[512] = "";
static thread_local int g_engine = 0;
int set_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}
void note_conv_engine(int engine) { g_engine = engine; }

// ---- reduction tickets (common.cuh) ----
static __device__ unsigned int g_tickets[kTicketSlots];        // zero at module load; every slot re-arms itself
int pdl_enabled() {
    static const int on = [] { const char* e = getenv("MSX_PDL"); return e ? atoi(e) : 1; }();
    return on;
}

unsigned int* next_ticket() {
    static std::atomic<unsigned> ctr{0};
    static std::mutex mu;
    static unsigned int* base[64] = {nullptr};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    unsigned int* b = base[dev];
    if (!b) {
        std::lock_guard<std::mutex> lk(mu);
        if (!base[dev]) {
            void* p = nullptr;
            if (cudaGetSymbolAddress(&p, g_tickets) != cudaSuccess) return nullptr;
            base[dev] = static_cast<unsigned int*>(p);
        }
        b = base[dev];
    }
    return b + ctr.fetch_add(1, std::memory_order_relaxed) % kTicketSlots;
}

// ---- profiler ----
struct ProfRec { cudaEvent_t a, b; int cls; double bytes, flops; };
static std::mutex g_mu;
static std::atomic<long long> g_launches{0};
static std::atomic<int> g_prof_on{0};
static std::vector<ProfRec*> g_recs;
static const char* kClassNames[K_NUM] = {
    "conv_gather_simt", "conv_wgrad_simt", "wgrad_reduce", "pack_weights", "bn_stats", "bn_finalize", "bn_apply",
    "bn_bwd_reduce", "bn_bwd_apply", "sum_n", "upsample2x_bwd", "maxpool", "layout", "head", "loss", "confmat", "optim",
    "conv_gather_tcgen05", "conv_wgrad_tcgen05", "misc"};

ProfScope::ProfScope(int c, cudaStream_t s, double bytes, double flops) : cls(c), st(s), rec(nullptr) {
    g_launches.fetch_add(1, std::memory_order_relaxed);
    if (!g_prof_on.load(std::memory_order_relaxed)) return;
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(s, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return;
    ProfRec* r = new ProfRec{nullptr, nullptr, c, bytes, flops};
    if (cudaEventCreate(&r->a) != cudaSuccess || cudaEventCreate(&r->b) != cudaSuccess) { delete r; return; }
    cudaEventRecord(r->a, s);
    rec = r;
}
ProfScope::~ProfScope() {
    if (!rec) return;
    ProfRec* r = static_cast<ProfRec*>(rec);
    cudaEventRecord(r->b, st);
    std::lock_guard<std::mutex> lk(g_mu);
    g_recs.push_back(r);
}
}  // namespace msx

extern "C" {
const char* msx_last_error(void) { return msx::g_err; }
int msx_version(void) { return 100; }
int msx_last_conv_engine(void) { return msx::g_engine; }

long long msx_launch_count(void) { return msx::g_launches.load(); }
int msx_profile_enable(int on) { msx::g_prof_on.store(on ? 1 : 0); return MSX_OK; }
int msx_profile_num_classes(void) { return msx::K_NUM; }
const char* msx_profile_class_name(int cls) { return (cls >= 0 && cls < msx::K_NUM) ? msx::kClassNames[cls] : "?"; }
/* Synchronises the device, folds all recorded launches into per-class totals and clears the records.
 * out arrays have msx_profile_num_classes() entries: milliseconds, launches, algorithmic bytes, flops. */
int msx_profile_collect(double* ms, long long* launches, double* bytes, double* flops) {
    using namespace msx;
    MSX_REQUIRE(ms && launches && bytes && flops, "profile_collect: null pointer");
    MSX_CHECK_CUDA(cudaDeviceSynchronize());
    std::lock_guard<std::mutex> lk(g_mu);
    for (int i = 0; i < K_NUM; ++i) { ms[i] = 0; launches[i] = 0; bytes[i] = 0; flops[i] = 0; }
    for (ProfRec* r : g_recs) {
        float t = 0.f;
        if (cudaEventElapsedTime(&t, r->a, r->b) == cudaSuccess) {
            ms[r->cls] += t; launches[r->cls] += 1; bytes[r->cls] += r->bytes; flops[r->cls] += r->flops;
        }
        cudaEventDestroy(r->a); cudaEventDestroy(r->b);
        delete r;
    }
    g_recs.clear();
    return MSX_OK;
}
}
