// bx_api.cu — error text, version and device queries of the C ABI (include/burgers_b200.h)
#include "bx_common.cuh"
#include <cstring>
#include <atomic>
#include <mutex>
#include <vector>

namespace bx {
static thread_local char g_err[512] = "";

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

// ---- tuning knobs (explicit API; nothing on the launch path reads the environment) -------------
static std::atomic<int> g_tune[BX_TUNE_COUNT];
static std::atomic<bool> g_tune_init{false};
static void tune_init() {
    bool expected = false;
    if (g_tune_init.compare_exchange_strong(expected, true))
        for (int i = 0; i < BX_TUNE_COUNT; ++i) g_tune[i].store(-1);
}
int tuning(int knob) {
    tune_init();
    return (knob >= 0 && knob < BX_TUNE_COUNT) ? g_tune[knob].load(std::memory_order_relaxed) : -1;
}

// ---- per-device properties, queried once ---------------------------------------------------------
int device_info(DeviceInfo &out) {
    static std::mutex mu;
    static DeviceInfo cache[64];
    static bool have[64] = {};
    int dev = 0;
    BX_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    if (dev < 0 || dev >= 64 || !have[dev]) {
        DeviceInfo d;
        d.device = dev;
        BX_CUDA(cudaDeviceGetAttribute(&d.sms, cudaDevAttrMultiProcessorCount, dev));
        BX_CUDA(cudaDeviceGetAttribute(&d.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
        BX_CUDA(cudaDeviceGetAttribute(&d.cooperative, cudaDevAttrCooperativeLaunch, dev));
        if (dev < 0 || dev >= 64) { out = d; return BX_OK; }
        cache[dev] = d; have[dev] = true;
    }
    out = cache[dev];
    return BX_OK;
}

// ---- kernel attributes: cudaFuncSetAttribute / occupancy once per (kernel, device, block, smem) ----
int kernel_occupancy(const void *kern, int threads, size_t smem, int *occ_out) {
    struct Key { const void *k; int dev, threads; size_t smem; int occ; };
    static std::mutex mu;
    static std::vector<Key> cache;
    int dev = 0;
    BX_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lock(mu);
    for (const Key &e : cache)
        if (e.k == kern && e.dev == dev && e.threads == threads && e.smem == smem) {
            if (occ_out) *occ_out = e.occ;
            return BX_OK;
        }
    // the opt-in size is ONE attribute per (function, device): only ever raise it, or a launch shape cached earlier
    // with a larger size would be launched against a lowered limit
    {
        size_t have = 0;
        for (const Key &e : cache)
            if (e.k == kern && e.dev == dev && e.smem > have) have = e.smem;
        if (smem > have)
            BX_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    }
    int occ = 0;
    BX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, threads, smem));
    cache.push_back(Key{kern, dev, threads, smem, occ});
    if (occ_out) *occ_out = occ;
    return BX_OK;
}

// ---- work-queue slots: every launch takes the next of the 64 counters of the caller's 256-byte
// workspace head, so launches that are in flight on different streams never share a counter.
unsigned int *queue_slot(void *workspace) {
    static std::atomic<unsigned> next{0};
    return static_cast<unsigned int *>(workspace) + (next.fetch_add(1, std::memory_order_relaxed) & 63u);
}
}  // namespace bx

extern "C" int bx_version(void) { return 200; }

extern "C" int bx_set_tuning(int knob, int value) {
    using namespace bx;
    BX_REQUIRE(knob >= 0 && knob < BX_TUNE_COUNT, "bx_set_tuning: unknown knob %d", knob);
    tune_init();
    g_tune[knob].store(value < 0 ? -1 : value);
    return BX_OK;
}

extern "C" int bx_get_tuning(int knob) { return bx::tuning(knob); }

extern "C" const char *bx_last_error(void) { return bx::g_err; }

extern "C" int bx_device_sm_count(void) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return BX_ERR_CUDA;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return BX_ERR_CUDA;
    return sms;
}
